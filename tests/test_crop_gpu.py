"""GPU parity: opb200_crop_warp (through the C ABI) vs the crop oracle and the cv2 fixtures -- bit-exact."""
import os

import numpy as np
import pytest
import torch

from openpose_b200 import crop_extractors as ce
from openpose_b200._lib import Opb200Error
from tests.cases_crop import CROP_CASES, crop_frame_input

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "crop_warp.npz")


@pytest.mark.parametrize("case", CROP_CASES, ids=[c[0] for c in CROP_CASES])
def test_crop_matches_cv2_fixture_and_oracle(port, cuda, case):
    name, seed, (hs, ws), size, m = case
    img = crop_frame_input(seed, hs, ws)
    got = ce.crop_frames_gpu(ce.upload_frame(img, cuda), [np.asarray(m)], size)[0].cpu().numpy()
    assert np.array_equal(got, ce.u_char_to_float(np.load(GOLDEN)[name]))
    assert np.array_equal(got, port.crop_warp(img, np.asarray(m), size))


def test_crop_batch_full_size(port, cuda):
    """1080p frame, 24 hand-sized crops (both mirrored and not, some hanging off the frame) in one launch."""
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, size=(1080, 1920, 3), dtype=np.uint8)
    mats = []
    for i in range(24):
        side = float(rng.uniform(40, 700))
        rect = (float(rng.uniform(-100, 1800)), float(rng.uniform(-100, 1000)), side, side)
        mats.append(ce.crop_matrix(rect, 368, mirror=bool(i & 1)))
    got = ce.crop_frames_gpu(ce.upload_frame(img, cuda), mats, (368, 368)).cpu().numpy()
    assert got.shape == (24, 3, 368, 368)
    for i in range(24):
        assert np.array_equal(got[i], port.crop_warp(img, mats[i], (368, 368))), i
    assert got.min() >= -0.5 and got.max() <= 255 / 256 - 0.5


def test_crop_errors(cuda):
    with pytest.raises(Opb200Error, match="Empty cvInputData"):
        ce.upload_frame(np.zeros((0, 0, 3), np.uint8), cuda)
    with pytest.raises(Opb200Error, match="3-channel"):
        ce.upload_frame(np.zeros((4, 4, 1), np.uint8), cuda)
