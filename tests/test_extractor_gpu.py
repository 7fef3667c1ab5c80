"""GPU tests of the extractor mirrors (PoseExtractorB200, HandExtractorB200, FaceExtractorB200) against the CPU
oracle chain on the same inputs.

* heat-maps given (enableNet = False): keypoints, scores, candidates and exported heat-maps are bit-exact against
  oracle resize -> reference nms -> reference connector (the north star's bit-exact clause);
* net on the path: multi-scale heat-maps within 2e-2 of the FP32 oracle (RefNet per scale -> oracle resize-merge);
* hand/face: the batched crop path equals crop-by-crop processing bit for bit, heat-maps are within 2e-2 of the
  FP32 oracle, and key-points are the back-projected first arg-max of each heat-map channel.
"""
import numpy as np
import pytest
import torch

from openpose_b200 import crop_extractors as ce
from openpose_b200 import pose_extractor as pe
from openpose_b200 import synth
from openpose_b200._lib import POSE_BODY_25
from oracle import ref_net
from tests.cases import make_case

pytestmark = pytest.mark.gpu
TOL_NET = 2e-2


def _oracle_pose(port, impl, net_out, in_size, scale_ratios=None, thr_props=None):
    """net_out: list of (C,h,w) -> (heat, peaks, keypoints, scores, scaleNetToOutput) per the reference chain."""
    h, w = net_out[0].shape[1] * 8, net_out[0].shape[2] * 8
    heat = port.resize_merge(net_out, h, w, scale_ratios)
    s = pe.resize_get_scale_factor(in_size, (w, h))
    net_size = (pe.positive_int_round(s * in_size[0]), pe.positive_int_round(s * in_size[1]))
    scale = float(np.float32(pe.resize_get_scale_factor(net_size, in_size)))
    off = float(np.float32(0.5 / scale))
    peaks = impl.nms(heat, 0.05, 25, (off, off))
    kp, sc = impl.connect(heat, peaks, 0, scale=scale)[:2]
    return heat, peaks, kp, sc, scale


@pytest.mark.parametrize("people", [0, 3, 12])
def test_pose_extractor_custom_heatmaps_bit_exact(port, cuda, people, request):
    impl = request.getfixturevalue("ref") if __import__("oracle.post", fromlist=["Ref"]).Ref.available() else port
    maps, _ = synth.planted_maps(0, 46, 82, people, seed=40 + people, sigma=1.2, paf_halfwidth=0.8, noise=0.005,
                                 limb_len=(3.0, 7.0))
    ex = pe.PoseExtractorB200(POSE_BODY_25, "unused", "unused", enableNet=False, addPartCandidates=True,
                              heatMapTypes=[pe.HeatMapType.Parts, pe.HeatMapType.Background, pe.HeatMapType.PAFs],
                              heatMapScaleMode=pe.ScaleMode.UnsignedChar)
    ex.initializationOnThread()
    x = np.zeros((1, 3, 368, 656), np.float32)           # only its shape matters when the net is disabled
    ex.forwardPass([x], (1280, 720), [1.0], maps[None])
    heat, peaks, kp, sc, scale = _oracle_pose(port, impl, [maps], (1280, 720))
    assert ex.getScaleNetToOutput() == scale
    assert ex.getHeatMapSize() == [1, 78, 368, 656]
    assert np.array_equal(ex.getHeatMapGpuConstPtr()[0].cpu().numpy(), heat)
    assert np.array_equal(ex.getPoseKeypoints(), kp) and np.array_equal(ex.getPoseScores(), sc)
    assert ex.getPoseKeypoints().shape[0] >= min(people, 1)
    cands = ex.getCandidatesCopy()
    assert len(cands) == 25
    for part in range(25):
        n = int(round(float(peaks[part, 0, 0])))
        want = [(float(np.float32(r[0]) * np.float32(scale)), float(np.float32(r[1]) * np.float32(scale)), float(r[2]))
                for r in peaks[part, 1:1 + n]]
        assert cands[part] == want
    hm = ex.getHeatMapsCopy()
    assert hm.shape == (78, 368, 656)
    f = np.float32
    assert np.array_equal(hm[:26], np.trunc(np.clip(heat[:26], 0, 1) * f(255.0) + f(0.5)))
    assert np.array_equal(hm[26:], np.trunc((np.clip(heat[26:], -1, 1) * f(128.5) + f(128.5)) + f(0.5)))
    # a second frame of another size reshapes every blob (poseExtractorCaffe.cpp:262-288)
    maps2, _ = make_case((0, 46, 82, 2, 0.01, 0.0, 18))
    ex.forwardPass([np.zeros((1, 3, 184, 328), np.float32)], (640, 360), [1.0], maps2[None, :, :23, :41])
    assert ex.getHeatMapSize() == [1, 78, 184, 328]


def test_pose_extractor_multiscale_net(port, cuda):
    """scale_number 2: two nets (one per scale, created on demand), merged heat-maps vs the FP32 oracle."""
    proto, cm, ref = ref_net.build("body_25", 1234)
    ex = pe.PoseExtractorB200(POSE_BODY_25, proto, cm, heatMapTypes=[pe.HeatMapType.Parts],
                              heatMapScaleMode=pe.ScaleMode.NoScale)
    ex.initializationOnThread()
    rng = np.random.default_rng(5)
    x0 = rng.uniform(-0.5, 0.5, (1, 3, 64, 96)).astype(np.float32)
    x1 = rng.uniform(-0.5, 0.5, (1, 3, 48, 72)).astype(np.float32)
    ex.forwardPass([x0, x1], (192, 128), [0.5, 0.375])
    assert len(ex.spNets) == 2 and ex.getHeatMapSize() == [1, 78, 64, 96]
    o0 = ref.forward(x0)["net_output"][0]
    o1 = ref.forward(x1)["net_output"][0]
    want = port.resize_merge([o0, o1], 64, 96, [float(np.float32(0.5)), float(np.float32(0.375))])
    got = ex.getHeatMapGpuConstPtr()[0].cpu().numpy()
    assert np.abs(got - want).max() <= TOL_NET
    assert np.array_equal(ex.getHeatMapsCopy(), got[:25])
    s = pe.resize_get_scale_factor((192, 128), (96, 64))
    assert ex.getScaleNetToOutput() == float(np.float32(pe.resize_get_scale_factor(
        (pe.positive_int_round(s * 192), pe.positive_int_round(s * 128)), (192, 128))))
    assert ex.getPoseKeypoints().shape[1:] == (25, 3) and ex.getPoseScores().shape[0] == ex.getPoseKeypoints().shape[0]


def _frame(seed, h=240, w=320):
    return np.random.default_rng(seed).integers(0, 256, (h, w, 3), dtype=np.uint8)


def test_hand_extractor_batched_equals_serial_and_oracle(port, cuda):
    proto, cm, ref = ref_net.build("hand", 1234)
    img = _frame(7)
    rects = [[(20.0, 30.0, 96.0, 96.0), (150.0, 60.0, 120.0, 120.0)],
             [(0.0, 0.0, 0.0, 0.0), (200.0, 100.0, 64.0, 64.0)]]          # person 1 has no left hand
    hx = ce.HandExtractorB200((368, 368), (368, 368), proto, cm, heatMapTypes=[pe.HeatMapType.Parts],
                              heatMapScaleMode=pe.ScaleMode.ZeroToOne)
    hx.initializationOnThread()
    hx.forwardPass(rects, img)
    kps = [k.copy() for k in hx.getHandKeypoints()]
    heat = [h.copy() for h in hx.getHeatMaps()]
    assert kps[0].shape == (2, 21, 3) and not kps[0][1].any()             # skipped rectangle stays zero
    # crop by crop (the reference's order of work) gives the same bits
    for hand, person in [(0, 0), (1, 0), (1, 1)]:
        single = [[(0.0, 0.0, 0.0, 0.0)] * 2]
        single[0][hand] = rects[person][hand]
        hx.forwardPass(single, img)
        assert np.array_equal(hx.getHandKeypoints()[hand][0], kps[hand][person])
        assert np.array_equal(hx.getHeatMaps()[hand][0], heat[hand][person])
    # one crop against the FP32 oracle: heat-maps within tolerance, key-points = back-projected arg-max
    crop = port.crop_warp(img, ce.crop_matrix(rects[0][1], 368, mirror=False), (368, 368))
    want = port.resize_merge([ref.forward(crop[None])["net_output"][0]], 368, 368)
    assert np.abs(np.clip(want[:21], 0, 1) - heat[1][0]).max() <= TOL_NET
    raw = hx._net.heat                                                    # last run: the (1, person 1) crop
    aff1 = ce.crop_matrix(rects[1][1], 368, mirror=False)
    peaks = port.maximum(raw[0].cpu().numpy(), 21)
    assert np.array_equal(ce.connect_keypoints(peaks, aff1), kps[1][1])


def test_hand_extractor_multiscale_keeps_best_scale(cuda):
    proto, cm, _ = ref_net.build("hand", 1234)
    img = _frame(8)
    rect = (60.0, 40.0, 100.0, 100.0)
    hx = ce.HandExtractorB200((368, 368), (368, 368), proto, cm, numberScales=3, rangeScales=0.4)
    hx.initializationOnThread()
    hx.forwardPass([[rect, (0.0, 0.0, 0.0, 0.0)]], img)
    got = hx.getHandKeypoints()[0][0].copy()
    single = ce.HandExtractorB200((368, 368), (368, 368), proto, cm)
    single.initializationOnThread()
    best = None
    for i in range(3):
        s = np.float32(0.8) + np.float32(0.4) * np.float32(i) / np.float32(2.0)
        side = float(pe.positive_int_round(float(np.float32(100.0) * s)) // 2 * 2)
        single.forwardPass([[ce.recenter(rect, side, side), (0.0, 0.0, 0.0, 0.0)]], img)
        est = single.getHandKeypoints()[0][0]
        if best is None or ce.get_average_score(est) > ce.get_average_score(best):
            best = est.copy()
    assert np.array_equal(got, best)


def test_face_extractor(port, cuda):
    proto, cm, ref = ref_net.build("face", 1234)
    img = _frame(9)
    rects = [(40.0, 30.0, 128.0, 128.0), (10.0, 10.0, 30.0, 30.0), (150.0, 80.0, 90.0, 90.0)]   # middle one too small
    fx = ce.FaceExtractorB200((368, 368), (368, 368), proto, cm, heatMapTypes=[pe.HeatMapType.Parts],
                              heatMapScaleMode=pe.ScaleMode.PlusMinusOne)
    fx.initializationOnThread()
    fx.forwardPass(rects, img)
    kp, heat = fx.getFaceKeypoints().copy(), fx.getHeatMaps().copy()
    assert kp.shape == (3, 70, 3) and not kp[1].any()
    raw = fx._net.heat.cpu().numpy()                     # crops of persons 0 and 2
    for j, person in enumerate((0, 2)):
        r = rects[person]
        m = np.array([[r[2] / 368.0, 0.0, r[0]], [0.0, r[2] / 368.0, r[1]]])
        assert np.array_equal(kp[person], ce.connect_keypoints(port.maximum(raw[j], 70), m))
        assert np.array_equal(heat[person], np.clip(raw[j][:70], 0, 1) * np.float32(2.0) - np.float32(1.0))
    m0 = np.array([[128.0 / 368.0, 0.0, 40.0], [0.0, 128.0 / 368.0, 30.0]])
    crop = port.crop_warp(img, m0, (368, 368))
    want = port.resize_merge([ref.forward(crop[None])["net_output"][0]], 368, 368)
    assert np.abs(want - raw[0]).max() <= TOL_NET
