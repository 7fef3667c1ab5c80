"""GPU: the reference's OWN hand and face extractors -- hand/handExtractorCaffe.cpp, handExtractorNet.cpp,
face/faceExtractorCaffe.cpp, faceExtractorNet.cpp, core/matrix.cpp and the *Caffe layer classes, compiled unmodified
(oracle/Makefile `cropextractor`) -- over the B200 backend. cv::Mat / cv::warpAffine come from the harness's functional
stand-in (warp = the cv2-pinned oracle restatement); driver tests/cpp/crop_main.cpp.

* keypoints and exported heat-maps equal the Python mirror (openpose_b200/crop_extractors.py: device crops, ONE batched
  forward for all crops) bit for bit -- the same library computes both, crop by crop there, batched here;
* one crop against the FP32 oracle chain (cv-pinned crop -> RefNet -> oracle resize): heat-maps within 2e-2;
* multi-scale hands keep the best-scoring scale (handExtractorCaffe.cpp:383-431)."""
import os
import shutil
import struct
import subprocess

import numpy as np
import pytest

from openpose_b200 import crop_extractors as ce
from openpose_b200 import pose_extractor as pe
from oracle import ref_net

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "crop_test")
TOL_NET = 2e-2


def _frame(seed, h=300, w=400):
    rng = np.random.default_rng(seed)
    f = rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8)
    f[:, : w // 3] = (np.arange(h)[:, None, None] * 3 % 256).astype(np.uint8)
    return f


def _model_folder(tmp_path, kind):
    """<folder>/hand/pose_deploy.prototxt + pose_iter_102000.caffemodel, or face/... (handParameters.hpp:44-45,
    faceParameters.hpp:25-26): the names the reference's constructors append to modelFolder."""
    proto, cm, ref = ref_net.build(kind, 1234)
    d = tmp_path / "models" / kind
    d.mkdir(parents=True, exist_ok=True)
    shutil.copy(proto, d / "pose_deploy.prototxt")
    shutil.copy(cm, d / ("pose_iter_102000.caffemodel" if kind == "hand" else "pose_iter_116000.caffemodel"))
    return str(tmp_path / "models") + "/", proto, cm, ref


def run_crop_extractor(tmp_path, kind, model_folder, frame, rects, scales=1, range_scales=0.4, types=0,
                       scale_mode=int(pe.ScaleMode.ZeroToOne)):
    if not os.path.exists(BIN):
        pytest.skip("oracle/_ref/crop_test not built (needs /root/reference at build time)")
    rects = np.ascontiguousarray(rects, np.float32)
    people = rects.shape[0]
    req = struct.pack("<7i", 0 if kind == "hand" else 1, people, scales, types, scale_mode, frame.shape[0], frame.shape[1])
    req += struct.pack("<f", range_scales)
    b = model_folder.encode()
    req += struct.pack("<i", len(b)) + b + np.ascontiguousarray(frame).tobytes() + rects.tobytes()
    path = tmp_path / "request.bin"
    path.write_bytes(req)
    prefix = str(tmp_path / "out")
    r = subprocess.run([BIN, str(path), prefix], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr

    def load(name):
        raw = open("%s.%s.bin" % (prefix, name), "rb").read()
        nd = struct.unpack_from("<i", raw)[0]
        dims = struct.unpack_from("<%di" % nd, raw, 4)
        return np.frombuffer(raw, np.float32, offset=4 + 4 * nd).reshape(dims).copy()

    names = ("left", "right", "heatleft", "heatright") if kind == "hand" else ("face", "heatface")
    return {k: load(k) for k in names}


def test_reference_hand_extractor(tmp_path, port, cuda):
    folder, proto, cm, ref = _model_folder(tmp_path, "hand")
    img = _frame(7)
    rects = [[(20.0, 30.0, 96.0, 96.0), (150.0, 60.0, 120.0, 120.0)],
             [(0.0, 0.0, 0.0, 0.0), (200.0, 100.0, 64.0, 64.0)]]          # person 1 has no left hand
    out = run_crop_extractor(tmp_path, "hand", folder, img, rects, types=1, scale_mode=int(pe.ScaleMode.ZeroToOne))
    assert out["left"].shape == (2, 21, 3) and not out["left"][1].any()
    hx = ce.HandExtractorB200((368, 368), (368, 368), proto, cm, heatMapTypes=[pe.HeatMapType.Parts],
                              heatMapScaleMode=pe.ScaleMode.ZeroToOne)
    hx.initializationOnThread()
    hx.forwardPass(rects, img)
    kps, heat = hx.getHandKeypoints(), hx.getHeatMaps()
    for hand, name in ((0, "left"), (1, "right")):
        assert np.array_equal(out[name], kps[hand])
    assert np.array_equal(out["heatleft"][0], heat[0][0])
    assert np.array_equal(out["heatright"], heat[1])
    crop = port.crop_warp(img, ce.crop_matrix(rects[0][1], 368, mirror=False), (368, 368))
    want = port.resize_merge([ref.forward(crop[None])["net_output"][0]], 368, 368)
    err = np.abs(np.clip(want[:21], 0, 1) - out["heatright"][0]).max()
    print("reference HandExtractorCaffe over the B200 backend: heat-maps vs FP32 oracle max-abs %.4f" % err)
    assert err <= TOL_NET


def test_reference_hand_extractor_multiscale(tmp_path, cuda):
    folder, proto, cm, _ = _model_folder(tmp_path, "hand")
    img = _frame(8)
    rect = (60.0, 40.0, 100.0, 100.0)
    out = run_crop_extractor(tmp_path, "hand", folder, img, [[rect, (0.0, 0.0, 0.0, 0.0)]], scales=3, range_scales=0.4)
    hx = ce.HandExtractorB200((368, 368), (368, 368), proto, cm, numberScales=3, rangeScales=0.4)
    hx.initializationOnThread()
    hx.forwardPass([[rect, (0.0, 0.0, 0.0, 0.0)]], img)
    assert np.array_equal(out["left"], hx.getHandKeypoints()[0])
    assert out["left"][0, :, 2].max() > 0


def test_reference_face_extractor(tmp_path, port, cuda):
    folder, proto, cm, ref = _model_folder(tmp_path, "face")
    img = _frame(9)
    rects = [(40.0, 30.0, 128.0, 128.0), (10.0, 10.0, 30.0, 30.0), (150.0, 80.0, 90.0, 90.0)]   # middle one too small
    out = run_crop_extractor(tmp_path, "face", folder, img, rects, types=1, scale_mode=int(pe.ScaleMode.PlusMinusOne))
    fx = ce.FaceExtractorB200((368, 368), (368, 368), proto, cm, heatMapTypes=[pe.HeatMapType.Parts],
                              heatMapScaleMode=pe.ScaleMode.PlusMinusOne)
    fx.initializationOnThread()
    fx.forwardPass(rects, img)
    assert out["face"].shape == (3, 70, 3) and not out["face"][1].any()
    assert np.array_equal(out["face"], fx.getFaceKeypoints())
    assert np.array_equal(out["heatface"][0], fx.getHeatMaps()[0]) and np.array_equal(out["heatface"][2], fx.getHeatMaps()[2])
    m0 = np.array([[128.0 / 368.0, 0.0, 40.0], [0.0, 128.0 / 368.0, 30.0]])
    crop = port.crop_warp(img, m0, (368, 368))
    want = port.resize_merge([ref.forward(crop[None])["net_output"][0]], 368, 368)
    got = (out["heatface"][0] + np.float32(1.0)) / np.float32(2.0)        # PlusMinusOne of the clamped maps
    err = np.abs(np.clip(want[:70], 0, 1) - got).max()
    print("reference FaceExtractorCaffe over the B200 backend: heat-maps vs FP32 oracle max-abs %.4f" % err)
    assert err <= TOL_NET
