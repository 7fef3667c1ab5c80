"""Shared synthetic cases for oracle and GPU parity tests (seeded, small enough for the CPU oracle)."""
import numpy as np

from openpose_b200 import synth

BODY_25, COCO_18 = 0, 1

# (model, h, w, people, noise, drop_prob, seed)
POST_CASES = [
    (BODY_25, 368, 656, 1, 0.01, 0.0, 11),
    (BODY_25, 368, 656, 8, 0.01, 0.1, 12),
    (BODY_25, 368, 656, 32, 0.01, 0.2, 13),
    (BODY_25, 368, 656, 100, 0.01, 0.0, 14),   # crowd (config 5)
    (BODY_25, 368, 656, 6, 0.03, 0.0, 15),     # noisy: every channel saturates at 127 peaks
    (BODY_25, 368, 656, 0, 0.005, 0.0, 16),    # empty frame
    (BODY_25, 184, 328, 5, 0.01, 0.3, 17),
    (BODY_25, 46, 82, 2, 0.01, 0.0, 18),       # tiny, odd sizes relative to 128-pixel chunks
    (BODY_25, 51, 67, 3, 0.02, 0.0, 19),       # h*w not a multiple of 4 -> scalar load path
    (COCO_18, 368, 656, 8, 0.01, 0.1, 21),
    (COCO_18, 368, 496, 40, 0.01, 0.0, 22),
    (COCO_18, 368, 656, 4, 0.03, 0.0, 23),
]


def make_case(case):
    model, h, w, people, noise, drop, seed = case
    sigma = 7.0 if h >= 184 else 1.5
    half = 4.0 if h >= 184 else 1.0
    limb = (18.0, 48.0) if h >= 184 else (4.0, 10.0)
    maps, planted = synth.planted_maps(model, h, w, people, seed, sigma=sigma, paf_halfwidth=half, noise=noise,
                                       drop_prob=drop, limb_len=limb)
    return maps, planted


def case_id(case):
    return "m%d_%dx%d_p%d_n%g_d%g" % case[:6]


def assert_peaks_equal(a, b):
    """Compare peak blobs up to each channel's count (rows beyond are stale by contract)."""
    assert a.shape == b.shape
    ca, cb = a[..., 0, 0].astype(int), b[..., 0, 0].astype(int)
    assert np.array_equal(ca, cb), (ca, cb)
    flat_a = a.reshape(-1, a.shape[-2], 3)
    flat_b = b.reshape(-1, b.shape[-2], 3)
    for c, n in enumerate(ca.reshape(-1)):
        assert np.array_equal(flat_a[c, : n + 1], flat_b[c, : n + 1]), "channel %d differs" % c
