"""The committed people fixture (tests/golden/people_golden.npz) still equals the reference's object code, plus the
host helper of KeypointScaler (getScaleAndOffset, keypointScaler.cpp:7-42) through the C ABI (no GPU needed)."""
import os

import numpy as np
import pytest

from oracle.post import Ref
from openpose_b200.people import KeypointScaler
from openpose_b200._lib import Opb200Error
from tests.cases_people import PEOPLE_CASES, SCALE_CASES, TOP_N, make_people

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "people_golden.npz")


@pytest.mark.skipif(not Ref.available(), reason="reference object code not built")
@pytest.mark.parametrize("case", PEOPLE_CASES, ids=[c[0] for c in PEOPLE_CASES])
def test_fixture_matches_reference_object_code(case):
    name, model, parts, people, seed = case
    ref, g = Ref(), np.load(GOLDEN)
    kp, sc = make_people(parts, people, seed)
    assert np.array_equal(g[name + "/hands"], ref.detect_hands(kp, model))
    assert np.array_equal(g[name + "/faces"], ref.detect_faces(kp, model))
    for n in TOP_N:
        assert np.array_equal(g["%s/top%d" % (name, n)], ref.keep_top_people(kp, sc, n))
    for i, (mode, s_io, s_no, prod) in enumerate(SCALE_CASES):
        assert np.array_equal(g["%s/scale%d" % (name, i)], ref.scale_keypoints(kp, mode, s_io, s_no, prod))


def test_fixture_covers_the_branches():
    g = np.load(GOLDEN)
    hands, faces = g["body25_12/hands"], g["body25_12/faces"]
    assert (hands[..., 2] > 0).any() and (hands[..., 2] == 0).any()
    assert (faces[:, 2] > 0).any() and (faces[:, 2] == 0).any()
    assert g["body25_40_ties/top5"].shape == (5, 25, 3) and g["body25_12/top50"].shape == (12, 25, 3)


def test_scale_and_offset_rules():
    f = np.float32
    assert KeypointScaler(2).scaleAndOffset(1.5, 2.0, (1280, 720)) == (0.0, 0.0, 1.5, 1.5)
    assert KeypointScaler(1).scaleAndOffset(1.5, 1.959, (1280, 720)) == (0.0, 0.0, float(f(1 / 1.959)), float(f(1 / 1.959)))
    assert KeypointScaler(3).scaleAndOffset(1, 1, (1280, 720)) == (0.0, 0.0, float(f(1) / f(1279)), float(f(1) / f(719)))
    assert KeypointScaler(4).scaleAndOffset(1, 1, (1280, 720)) == (0.0, 0.0, float(f(1) / f(1279)), float(f(1) / f(1279)))
    assert KeypointScaler(5).scaleAndOffset(1, 1, (1280, 720)) == (-1.0, -1.0, float(f(2) / f(1279)), float(f(2) / f(719)))
    assert KeypointScaler(6).scaleAndOffset(1, 1, (333, 777)) == (-1.0, -1.0, float(f(2) / f(776)), float(f(2) / f(776)))
    assert KeypointScaler(0).scaleAndOffset(9, 9, (10, 10)) == (0.0, 0.0, 1.0, 1.0)
    with pytest.raises(Opb200Error, match="Unknown ScaleMode selected."):
        KeypointScaler(8).scaleAndOffset(1, 1, (10, 10))
