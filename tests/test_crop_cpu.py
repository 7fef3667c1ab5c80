"""Pins the crop oracle (oracle/post_oracle.c: opo_crop_warp) on OpenCV: the committed cv2.warpAffine fixtures and,
when cv2 is importable, cv2 itself on further random cases."""
import os

import numpy as np
import pytest

from openpose_b200 import crop_extractors as ce
from tests.cases_crop import CROP_CASES, crop_frame_input

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "crop_warp.npz")


def _to_net_input(img_hwc):
    return ce.u_char_to_float(img_hwc)


@pytest.mark.parametrize("case", CROP_CASES, ids=[c[0] for c in CROP_CASES])
def test_oracle_matches_cv2_fixture(port, case):
    name, seed, (hs, ws), size, m = case
    g = np.load(GOLDEN)
    img = crop_frame_input(seed, hs, ws)
    assert g[name].shape == (size[1], size[0], 3)
    assert np.array_equal(port.crop_warp(img, np.asarray(m), size), _to_net_input(g[name]))


def test_oracle_matches_cv2_live(port):
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(77)
    img = rng.integers(0, 256, size=(270, 480, 3), dtype=np.uint8)
    for _ in range(12):
        sc = float(rng.uniform(0.05, 3.0))
        mirror = bool(rng.integers(0, 2))
        m = np.array([[-sc if mirror else sc, 0.0, float(np.float32(rng.uniform(-50, 500)))],
                      [0.0, sc, float(np.float32(rng.uniform(-50, 300)))]])
        ref = cv2.warpAffine(img, m, (184, 184), flags=cv2.INTER_LINEAR | cv2.WARP_INVERSE_MAP,
                             borderMode=cv2.BORDER_CONSTANT, borderValue=(0, 0, 0))
        assert np.array_equal(port.crop_warp(img, m, (184, 184)), _to_net_input(ref))


def test_crop_matrix_is_the_reference_float_arithmetic():
    # scale = rect.width / side in float; mirrored offset = float(x + width)
    m = ce.crop_matrix((10.3, 20.7, 100.1, 100.1), 368, mirror=True)
    sc = np.float32(100.1) / np.float32(368)
    assert m[0, 0] == -float(sc) and m[1, 1] == float(sc)
    assert m[0, 2] == float(np.float32(10.3) + np.float32(100.1)) and m[1, 2] == float(np.float32(20.7))
    assert m[0, 1] == 0 and m[1, 0] == 0
