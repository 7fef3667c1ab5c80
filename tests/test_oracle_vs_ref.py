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
