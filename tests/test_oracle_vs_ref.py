"""Pins the C restatement (oracle/post_oracle.c) against the reference's own object code
(oracle/_ref/libopref_post.so, built from /root/reference sources) -- bit-exact."""
import numpy as np
import pytest

from oracle.post import BODY_25, COCO_18, model_dims, pose_defaults
from openpose_b200 import synth
from tests.cases import POST_CASES, assert_peaks_equal, case_id, make_case


@pytest.mark.parametrize("model", [BODY_25, COCO_18])
def test_model_tables_and_defaults(ref, model):
    pairs, mapidx = ref.tables(model)
    parts, p, m = synth.model_tables(model)
    assert pairs == p and mapidx == m
    assert (parts, len(p) // 2) == model_dims(model)
    for mp in (False, True):
        a, b = ref.defaults(model, mp), pose_defaults(model, mp)
        for k in a:
            assert np.float32(a[k]) == np.float32(b[k]), k


@pytest.mark.parametrize("case", POST_CASES, ids=case_id)
def test_nms_and_connect_match_reference(port, ref, case):
    model = case[0]
    parts = model_dims(model)[0]
    maps, _ = make_case(case)
    pr = ref.nms(maps, 0.05, parts, off=(0.25, 0.25))
    pp = port.nms(maps, 0.05, parts, off=(0.25, 0.25))
    assert_peaks_equal(pr, pp)
    for mp in (False, True):
        kr, sr = ref.connect(maps, pr, model, scale=1.959, maximize_positives=mp)
        kp, sp, rows = port.connect(maps, pr, model, scale=1.959, maximize_positives=mp)
        assert kr.shape == kp.shape
        assert np.array_equal(kr, kp) and np.array_equal(sr, sp)
        # CUDA-build host half (pafPtrIntoVector + pafVectorIntoPeopleVector) on the same score cube
        cube = port.pair_scores(maps, pr, model)
        kr2, sr2 = ref.connect_from_pair_scores(cube, pr, model, scale=1.959, maximize_positives=mp)
        kp2, sp2, _ = port.connect_from_pair_scores(cube, pr, model, scale=1.959, maximize_positives=mp)
        assert kr2.shape == kp2.shape
        assert np.array_equal(kr2, kp2) and np.array_equal(sr2, sp2)


def test_nms_threshold_validation(port, ref):
    maps = np.zeros((26, 16, 16), np.float32)
    with pytest.raises(ValueError):
        port.nms(maps, 1.5, 25)
    with pytest.raises(ValueError):
        ref.nms(maps, 1.5, 25)


def test_nms_ring_rule_and_ties(port, ref):
    """Plateaus, ring (first inner border) '>=' rule and outer-border pixels, nmsBase.cpp:41-67."""
    rng = np.random.default_rng(5)
    for h, w in [(8, 9), (5, 5), (4, 7), (12, 6), (3, 3), (2, 5)]:
        for _ in range(20):
            m = (rng.integers(0, 4, size=(3, h, w)) / 4.0).astype(np.float32)  # many exact ties
            a = ref.nms(m, 0.05, 2)
            b = port.nms(m, 0.05, 2)
            assert_peaks_equal(a, b)


def test_planted_people_are_recovered(port):
    """Sanity of the generator + oracle: a clean planted frame yields the planted number of people."""
    maps, planted = synth.planted_maps(BODY_25, 368, 656, 6, seed=3, noise=0.0)
    peaks = port.nms(maps, 0.05, 25)
    kp, sc, rows = port.connect(maps, peaks, BODY_25)
    assert kp.shape[0] == 6
    # every planted part is found within 1 px
    for p in range(6):
        d = np.abs(kp[:, :, :2][None] - planted[p, :, :2][None, None]).max(-1)  # (1, P, parts)
        assert (d.min(axis=1) < 1.0).all()


# ---- bicubic arithmetic: the C restatement against cuda.hu itself, compiled as host C++ into oracle/_ref ----
def _ulp_diff(a, b):
    """max |a - b| in units of the float32 spacing at the larger magnitude of the two."""
    return float((np.abs(a.astype(np.float64) - b) / np.spacing(np.maximum(np.abs(a), np.abs(b)).astype(np.float32)
                                                                 + np.float32(1e-30))).max())


@pytest.mark.parametrize("shape", [(3, 46, 82), (2, 5, 7), (2, 1, 1), (1, 23, 31)])
def test_resize8_as_written_equals_cuda_hu(port, ref, shape):
    """contract=False == resize8TimesKernel's walk over cuda.hu's own bicubicInterpolate / bicubicInterpolate8Times
    (incl. its 5x5 shared tile and the first-block special case), bit for bit."""
    rng = np.random.default_rng(1)
    src = rng.normal(0, 1, size=shape).astype(np.float32)
    c, h, w = shape
    want = ref.hu_resize8(src)
    got = port.resize_merge([src], 8 * h, 8 * w, contract=False)
    assert np.array_equal(want, got)
    # the CUDA build's FMA-contracted evaluation (what the GPU tests pin on the reference's kernels) stays within
    # a few units of the last place of the output's scale
    fused = port.resize_merge([src], 8 * h, 8 * w, contract=True)
    scale = np.abs(src).max()
    assert np.abs(fused - want).max() <= 8 * np.spacing(np.float32(4 * scale)), np.abs(fused - want).max()
    assert not np.array_equal(fused, want) or h * w == 1


def test_resize_multiscale_as_written_equals_cuda_hu(port, ref):
    rng = np.random.default_rng(2)
    sizes = [(46, 82), (34, 60), (22, 40), (10, 20)]
    ratios = [1.0, 0.75, 0.5, 0.25]
    srcs = [rng.normal(0, 1, size=(3, h, w)).astype(np.float32) for h, w in sizes]
    want = ref.hu_resize_average(srcs, 368, 656, ratios)
    got = port.resize_merge(srcs, 368, 656, ratios, contract=False)
    assert np.array_equal(want, got)
    fused = port.resize_merge(srcs, 368, 656, ratios, contract=True)
    assert np.abs(fused - want).max() <= 8 * np.spacing(np.float32(4 * max(np.abs(s).max() for s in srcs)))


@pytest.mark.parametrize("hs,ws,net,scale", [(20, 31, (24, 40), 1.0), (48, 64, (32, 48), 0.6), (9, 13, (16, 16), 1.2307693)])
def test_preprocess_as_written_equals_cuda_hu(port, ref, hs, ws, net, scale):
    rng = np.random.default_rng(hs)
    frame = rng.integers(0, 256, size=(hs, ws, 3), dtype=np.uint8)
    planes = frame.astype(np.float32).transpose(2, 0, 1) * np.float32(1 / 256.0) - np.float32(0.5)  # cuda.cu:23, exact
    want = ref.hu_resize_and_pad(planes, net, scale)
    assert np.array_equal(want, port.preprocess(frame, net, scale, contract=False))
    assert np.abs(port.preprocess(frame, net, scale, contract=True) - want).max() <= 8 * np.spacing(np.float32(2))


@pytest.mark.parametrize("shape", [(22, 46, 46), (4, 5, 7), (3, 33, 31)])
def test_maximum_equals_cv_min_max_loc(port, shape):
    """maximumCpu is cv::minMaxLoc per channel (maximumBase.cpp:28-52); cv2 is importable here."""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(3)
    c, h, w = shape
    src = rng.normal(0, 1, size=(c, h, w)).astype(np.float32)
    src[0] = np.round(src[0])       # ties: both return the first maximum in raster order
    src[1] = 0.25
    got = port.maximum(src, c - 1)
    for ch in range(c - 1):
        _, vmax, _, loc = cv2.minMaxLoc(src[ch])
        assert (got[ch, 0], got[ch, 1], got[ch, 2]) == (np.float32(loc[0]), np.float32(loc[1]), np.float32(vmax))
