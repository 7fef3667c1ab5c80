"""The C restatement in its CUDA-build mode (FMA-contracted bicubic, nmsGpu's border rule and centroid, maximumGpu, the
device pre-processing) against golden vectors produced by the reference's OWN CUDA kernels on a B200
(tests/golden/refcuda_golden.npz, generator tests/golden/make_refcuda_golden.py). Runs on the CPU, without /root/reference,
without oracle/_ref and without a GPU: the oracle is pinned on the reference's GPU code in the CPU suite too."""
import os
import zlib

import numpy as np
import pytest

from oracle.post import MODE_CUDA
from tests import cases_refcuda as K
from tests.cases import assert_peaks_equal

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "refcuda_golden.npz")
pytestmark = pytest.mark.skipif(not os.path.exists(GOLDEN), reason="refcuda golden vectors not generated yet")


@pytest.fixture(scope="module")
def golden():
    return np.load(GOLDEN)


def _crc(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes())


@pytest.mark.parametrize("case", K.RESIZE_CASES, ids=[c[0] for c in K.RESIZE_CASES])
def test_resize_contracted_matches_reference_kernels(port, golden, case):
    name, sizes, ratios, target, planes, seed = case
    srcs = K.resize_inputs(sizes, planes, seed)
    assert _crc(np.concatenate([s.ravel() for s in srcs])) == int(golden["resize/%s/crc" % name][0]), "generator drifted"
    got = port.resize_merge([s[0] for s in srcs], target[0], target[1], ratios, contract=True)
    assert np.array_equal(got, golden["resize/%s/out" % name])


@pytest.mark.parametrize("case", K.NMS_CASES, ids=[c[0] for c in K.NMS_CASES])
def test_nms_cuda_mode_matches_reference_kernels(port, golden, case):
    name, parts, h, w, seed = case
    heat = K.nms_input(parts, h, w, seed)
    assert _crc(heat) == int(golden["nms/%s/crc" % name][0]), "generator drifted"
    assert_peaks_equal(golden["nms/%s/peaks" % name], port.nms(heat, 0.05, parts, off=(0.5, 0.5), mode=MODE_CUDA))


@pytest.mark.parametrize("case", K.MAX_CASES, ids=[c[0] for c in K.MAX_CASES])
def test_maximum_matches_reference_kernels(port, golden, case):
    name, c, h, w, seed = case
    heat = K.max_input(c, h, w, seed)
    assert _crc(heat) == int(golden["max/%s/crc" % name][0]), "generator drifted"
    assert np.array_equal(port.maximum(heat, c - 1), golden["max/%s/out" % name])


@pytest.mark.parametrize("case", K.PRE_CASES, ids=[c[0] for c in K.PRE_CASES])
def test_preprocess_contracted_matches_reference_kernels(port, golden, case):
    name, hs, ws, net, scale, seed = case
    frame = K.pre_input(hs, ws, seed)
    assert _crc(frame) == int(golden["pre/%s/crc" % name][0]), "generator drifted"
    assert np.array_equal(port.preprocess(frame, net, scale, contract=True), golden["pre/%s/out" % name])


@pytest.mark.parametrize("ci", K.CONNECT_CASES)
def test_connect_cuda_semantics_matches_reference_kernels(port, golden, ci):
    """pafScoreKernel's cube (process() with nvcc's FMA contraction) and connectBodyPartsGpu's people."""
    from oracle.post import model_dims, pose_defaults
    from tests.cases import POST_CASES, case_id, make_case
    case = POST_CASES[ci]
    model = case[0]
    parts = model_dims(model)[0]
    maps, _ = make_case(case)
    cid = case_id(case)
    if "connect/%s/crc" % cid not in golden:
        pytest.skip("fixture generated before the connector cases were added")
    assert _crc(maps) == int(golden["connect/%s/crc" % cid][0]), "generator drifted"
    peaks = port.nms(maps, 0.05, parts, off=(0.25, 0.25), mode=MODE_CUDA)
    for mp in (0, 1):
        cube = port.pair_scores(maps, peaks, model, p=pose_defaults(model, bool(mp)), mode=MODE_CUDA)
        assert np.array_equal(cube, golden["connect/%s/cube%d" % (cid, mp)])
        res = port.connect_from_pair_scores(cube, peaks, model, scale=1.959, maximize_positives=bool(mp))
        assert np.array_equal(res[0], golden["connect/%s/kp%d" % (cid, mp)])
        assert np.array_equal(res[1], golden["connect/%s/sc%d" % (cid, mp)])
