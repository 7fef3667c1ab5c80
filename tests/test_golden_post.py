"""The C restatement against the committed golden vectors (generated from the reference's object code by
tests/golden/make_golden.py). Runs without /root/reference and without oracle/_ref."""
import os
import zlib

import numpy as np
import pytest

from oracle.post import model_dims
from tests.cases import POST_CASES, assert_peaks_equal, case_id, make_case

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "post_golden.npz")


@pytest.fixture(scope="module")
def golden():
    return np.load(GOLDEN)


@pytest.mark.parametrize("case", POST_CASES, ids=case_id)
def test_port_matches_golden(port, golden, case):
    model = case[0]
    parts = model_dims(model)[0]
    maps, _ = make_case(case)
    cid = case_id(case)
    assert zlib.crc32(maps.tobytes()) == int(golden[cid + "/crc"][0]), "synthetic generator drifted"
    peaks = port.nms(maps, 0.05, parts, off=(0.25, 0.25))
    assert_peaks_equal(golden[cid + "/peaks"], peaks)
    for mp in (0, 1):
        kp, sc, _ = port.connect(maps, peaks, model, scale=1.959, maximize_positives=bool(mp))
        assert np.array_equal(kp, golden["%s/kp%d" % (cid, mp)])
        assert np.array_equal(sc, golden["%s/sc%d" % (cid, mp)])
