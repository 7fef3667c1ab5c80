"""GPU: the reference's OWN body extractor -- pose/poseExtractorCaffe.cpp, pose/poseExtractorNet.cpp and the four
net/*Caffe.cpp layer classes, compiled unmodified from /root/reference with -DUSE_CAFFE -DUSE_CUDA (oracle/Makefile
`extractor`) -- running over the B200 backend (integration/netCaffeB200.cpp = op::NetCaffe, postB200.cpp = the *Gpu
free functions, arrayCpuGpuB200.cpp = op::ArrayCpuGpu, libopb200.so). The driver tests/cpp/extractor_main.cpp calls
PoseExtractorCaffe::forwardPass + the getters exactly as WPoseExtractor does and dumps every output.

* custom net output (planted maps): keypoints / scores / peaks / candidates / heat-maps bit-exact against the oracle
  chain, in BOTH semantics -- CPU_ONLY build (reference object code: nmsCpu + connectBodyPartsCpu) and CUDA build
  (the reference's own kernels, oracle/_ref/libopref_cuda.so, on the same GPU);
* net on the path (1 and 2 scales, BODY_25 and COCO): the heat-map blob within 2e-2 of the FP32 oracle, and the rest
  of the chain bit-exact on the extractor's own heat-maps."""
import os
import struct
import subprocess

import numpy as np
import pytest
import torch

from openpose_b200 import pose_extractor as pe
from openpose_b200 import synth
from oracle import ref_net
from oracle.post import Ref, pose_defaults

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "extractor_test")
PARTS, BKG, PAFS = 1, 2, 4


def _arr(a):
    a = np.ascontiguousarray(a, np.float32)
    assert a.ndim == 4
    return struct.pack("<4i", *a.shape) + a.tobytes()


def _string(s):
    b = s.encode()
    return struct.pack("<i", len(b)) + b


def run_extractor(tmp_path, model, inputs, scales, in_size, net_output=None, proto="", weights="", types=0,
                  scale_mode=int(pe.ScaleMode.NoScale), candidates=True, maximize=False, repeats=1, semantics="cpu"):
    if not os.path.exists(BIN):
        pytest.skip("oracle/_ref/extractor_test not built (needs /root/reference at build time)")
    req = struct.pack("<10i", model, int(net_output is None), len(inputs), in_size[0], in_size[1], types, scale_mode,
                      int(candidates), int(maximize), repeats)
    req += _string("") + _string(proto) + _string(weights)   # modelFolder + path (poseExtractorCaffe.cpp:78-81)
    for s, x in zip(scales, inputs):
        req += struct.pack("<d", s) + _arr(x)
    if net_output is not None:
        req += _arr(net_output)
    path = tmp_path / "request.bin"
    path.write_bytes(req)
    prefix = str(tmp_path / "out")
    env = dict(os.environ, OPB200_SEMANTICS=semantics)
    r = subprocess.run([BIN, str(path), prefix], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout + r.stderr

    def load(name):
        raw = open("%s.%s.bin" % (prefix, name), "rb").read()
        nd = struct.unpack_from("<i", raw)[0]
        dims = struct.unpack_from("<%di" % nd, raw, 4)
        return np.frombuffer(raw, np.float32, offset=4 + 4 * nd).reshape(dims).copy()

    out = {k: load(k) for k in ("keypoints", "scores", "heatblob", "peaks", "heatmaps", "candidates", "scale")}
    out["stdout"] = r.stdout
    return out


def _scale_net_to_output(heat_hw, in_size):
    """poseExtractorCaffe.cpp:306-310."""
    h, w = heat_hw
    s = pe.resize_get_scale_factor(in_size, (w, h))
    net_size = (pe.positive_int_round(s * in_size[0]), pe.positive_int_round(s * in_size[1]))
    return float(np.float32(pe.resize_get_scale_factor(net_size, in_size)))


def _check_candidates(flat, peaks, scale, parts):
    """getCandidatesCopy (poseExtractorNet.cpp:255-273): peaks scaled by mScaleNetToOutput."""
    assert int(flat[0]) == parts
    pos = 1
    for p in range(parts):
        n = int(flat[pos])
        pos += 1
        assert n == int(round(float(peaks[p, 0, 0])))
        got = flat[pos:pos + 3 * n].reshape(n, 3)
        pos += 3 * n
        want = peaks[p, 1:1 + n].copy()
        want[:, 0] = want[:, 0] * np.float32(scale)
        want[:, 1] = want[:, 1] * np.float32(scale)
        assert np.array_equal(got, want)
    assert pos == flat.size


def _check_chain_cpu(out, heat, model, parts, in_size, port, maximize=False):
    """peaks / keypoints / scores of the extractor == reference nmsCpu + connectBodyPartsCpu on `heat`."""
    impl = Ref() if Ref.available() else port
    scale = _scale_net_to_output(heat.shape[1:], in_size)
    assert float(out["scale"][0]) == scale
    off = float(np.float32(0.5 / scale))
    d = pose_defaults(model, maximize)
    peaks = impl.nms(heat, d["nms_threshold"], parts, (off, off))
    for c in range(parts):
        n = int(peaks[c, 0, 0])
        assert n == int(out["peaks"][c, 0, 0])
        assert np.array_equal(peaks[c, 1:n + 1], out["peaks"][c, 1:n + 1])
    kp, sc = impl.connect(heat, peaks, model, scale=scale, maximize_positives=maximize)[:2]
    if kp.shape[0] == 0:
        assert out["keypoints"].size == 0 and out["scores"].size == 0
    else:
        assert np.array_equal(out["keypoints"], kp) and np.array_equal(out["scores"], sc)
    _check_candidates(out["candidates"], peaks, scale, parts)
    return kp.shape[0]


@pytest.mark.parametrize("people", [0, 3, 12])
def test_reference_extractor_custom_net_output_cpu_semantics(tmp_path, port, cuda, people):
    maps, _ = synth.planted_maps(0, 46, 82, people, seed=40 + people, sigma=1.2, paf_halfwidth=0.8, noise=0.005,
                                 limb_len=(3.0, 7.0))
    x = np.zeros((1, 3, 368, 656), np.float32)      # only its shape matters when the net is disabled
    out = run_extractor(tmp_path, 0, [x], [1.0], (1280, 720), net_output=maps[None], types=PARTS | BKG | PAFS,
                        scale_mode=int(pe.ScaleMode.UnsignedChar))
    heat = port.resize_merge([maps], 368, 656)
    assert out["heatblob"].shape == (1, 78, 368, 656)
    assert np.array_equal(out["heatblob"][0], heat)
    n = _check_chain_cpu(out, heat, 0, 25, (1280, 720), port)
    assert n >= min(people, 1)
    assert ("extractor ok: %d people" % n) in out["stdout"]
    f = np.float32
    hm = out["heatmaps"]
    assert hm.shape == (78, 368, 656)
    assert np.array_equal(hm[:26], np.trunc(np.clip(heat[:26], 0, 1) * f(255.0) + f(0.5)))
    assert np.array_equal(hm[26:], np.trunc((np.clip(heat[26:], -1, 1) * f(128.5) + f(128.5)) + f(0.5)))


@pytest.mark.parametrize("people", [0, 5, 30])
def test_reference_extractor_custom_net_output_cuda_semantics(tmp_path, cuda, people):
    """OPB200_SEMANTICS=cuda: the reference's extractor over the B200 backend == the reference's CUDA kernels."""
    from oracle.ref_cuda import RefCuda, REFCUDA_SO
    if not os.path.exists(REFCUDA_SO):
        pytest.skip("oracle/_ref/libopref_cuda.so not built")
    rc = RefCuda()
    maps, _ = synth.planted_maps(0, 46, 82, people, seed=70 + people, sigma=1.2, paf_halfwidth=0.8, noise=0.005,
                                 limb_len=(3.0, 7.0))
    x = np.zeros((1, 3, 368, 656), np.float32)
    out = run_extractor(tmp_path, 0, [x], [1.0], (1280, 720), net_output=maps[None], semantics="cuda")
    src = torch.from_numpy(maps[None]).to(cuda)
    heat = torch.empty((1, 78, 368, 656), device=cuda)
    rc.resize_merge(heat, [src])
    assert np.array_equal(out["heatblob"], heat.cpu().numpy())
    scale = _scale_net_to_output((368, 656), (1280, 720))
    off = float(np.float32(0.5 / scale))
    peaks = torch.zeros((1, 25, 128, 3), device=cuda)
    rc.nms(peaks, heat, 0.05, (off, off))
    pk = peaks[0].cpu().numpy()
    for c in range(25):
        n = int(pk[c, 0, 0])
        assert n == int(out["peaks"][c, 0, 0])
        assert np.array_equal(pk[c, 1:n + 1], out["peaks"][c, 1:n + 1])
    kp, sc, _ = rc.connect(heat[0], peaks[0], 0, scale=scale, want_cube=False)
    if kp.shape[0] == 0:
        assert out["keypoints"].size == 0
    else:
        assert np.array_equal(out["keypoints"], kp) and np.array_equal(out["scores"], sc)
    assert kp.shape[0] >= min(people, 1)


@pytest.mark.parametrize("model,net,in_size", [("body_25", (368, 656), (1280, 720)), ("coco", (184, 328), (640, 360))])
def test_reference_extractor_with_net_single_scale(tmp_path, port, cuda, model, net, in_size):
    proto, cm, ref = ref_net.build(model, 1234)
    mid = 0 if model == "body_25" else 1
    parts, ch = (25, 78) if mid == 0 else (18, 57)
    h, w = net
    x = synth.frames_to_net_input(synth.frames(1, h, w, 9))
    out = run_extractor(tmp_path, mid, [x], [1.0], in_size, proto=proto, weights=cm, types=PARTS, repeats=2)
    assert out["heatblob"].shape == (1, ch, h, w)
    want = port.resize_merge([ref.forward(x, keep=["net_output"])["net_output"][0]], h, w)
    err = np.abs(out["heatblob"][0] - want).max()
    print("%s extractor heat-maps vs FP32 oracle: max-abs %.3g (rms %.3g)" % (model, err, np.sqrt((want ** 2).mean())))
    assert err <= 2e-2
    _check_chain_cpu(out, out["heatblob"][0], mid, parts, in_size, port)
    assert np.array_equal(out["heatmaps"], out["heatblob"][0, :parts])     # NoScale, Parts only


def test_reference_extractor_with_net_two_scales(tmp_path, port, cuda):
    """scale_number 2: a second NetCaffe is created on demand (poseExtractorCaffe.cpp:241-245), the merge averages."""
    proto, cm, ref = ref_net.build("body_25", 1234)
    x0 = synth.frames_to_net_input(synth.frames(1, 128, 192, 3))
    x1 = synth.frames_to_net_input(synth.frames(1, 96, 144, 4))
    s0, s1 = 0.5, 0.375
    out = run_extractor(tmp_path, 0, [x0, x1], [s0, s1], (384, 256), proto=proto, weights=cm)
    o0 = ref.forward(x0, keep=["net_output"])["net_output"][0]
    o1 = ref.forward(x1, keep=["net_output"])["net_output"][0]
    want = port.resize_merge([o0, o1], 128, 192, [float(np.float32(s0)), float(np.float32(s1))])
    assert out["heatblob"].shape == (1, 78, 128, 192)
    assert np.abs(out["heatblob"][0] - want).max() <= 2e-2
    _check_chain_cpu(out, out["heatblob"][0], 0, 25, (384, 256), port)


@pytest.mark.parametrize("scale_mode", [pe.ScaleMode.ZeroToOne, pe.ScaleMode.PlusMinusOne, pe.ScaleMode.UnsignedChar,
                                        pe.ScaleMode.NoScale, pe.ScaleMode.ZeroToOneFixedAspect,
                                        pe.ScaleMode.PlusMinusOneFixedAspect])
@pytest.mark.parametrize("types", [PARTS, PARTS | BKG | PAFS, BKG | PAFS])
def test_heatmap_export_kernel_equals_reference_get_heat_maps_copy(tmp_path, port, cuda, scale_mode, types):
    """f2: opb200_heatmaps_export (one kernel) == the reference's own PoseExtractorNet::getHeatMapsCopy
    (poseExtractorNet.cpp:106-243, compiled unmodified into the harness) on the same heat-map blob, bit for bit."""
    from openpose_b200 import post
    maps, _ = synth.planted_maps(0, 46, 82, 4, seed=321, sigma=1.2, paf_halfwidth=0.8, noise=0.02, limb_len=(3.0, 7.0))
    maps[3] *= 1.7                                    # values beyond [0, 1] / [-1, 1]: the clamps matter
    maps[40] *= -2.5
    x = np.zeros((1, 3, 368, 656), np.float32)
    out = run_extractor(tmp_path, 0, [x], [1.0], (1280, 720), net_output=maps[None], types=types,
                        scale_mode=int(scale_mode))
    heat = torch.from_numpy(out["heatblob"]).to(cuda)
    got = post.heatmaps_export_gpu(heat, 25, True, 52, types, int(scale_mode), to_host=False)[0].cpu().numpy()
    assert got.shape == out["heatmaps"].shape
    assert np.array_equal(got, out["heatmaps"])
