"""Seeded inputs for the hand / face crop (cv::warpAffine INTER_LINEAR | WARP_INVERSE_MAP) parity tests."""
import numpy as np

# name, seed, (Hs, Ws), cv::Size (Wt, Ht), 2x3 inverse map
CROP_CASES = [
    ("hand_right", 1, (120, 160), (92, 92), [[0.4130434, 0.0, 30.0], [0.0, 0.4130434, 20.0]]),
    ("hand_left_mirror", 2, (120, 160), (92, 92), [[-0.6521739, 0.0, 110.0], [0.0, 0.6521739, 35.0]]),
    ("off_frame_corner", 3, (90, 70), (64, 48), [[1.37, 0.0, -20.5], [0.0, 1.37, 60.25]]),
    ("identity", 4, (50, 60), (60, 50), [[1.0, 0.0, 0.0], [0.0, 1.0, 0.0]]),
    ("magnify", 5, (40, 40), (96, 96), [[0.0271739, 0.0, 10.5], [0.0, 0.0271739, 20.25]]),
    ("half_pixel_ties", 6, (33, 47), (40, 30), [[0.5, 0.0, 0.015625], [0.0, 0.5, 0.484375]]),
    ("sheared", 7, (100, 120), (80, 60), [[0.8, 0.3, 10.2], [-0.25, 0.9, 20.7]]),
    ("face", 8, (180, 240), (92, 92), [[1.3913043, 0.0, 40.0], [0.0, 1.3913043, 30.0]]),
]


def crop_frame_input(seed: int, hs: int, ws: int) -> np.ndarray:
    rng = np.random.default_rng(1000 + seed)
    img = rng.integers(0, 256, size=(hs, ws, 3), dtype=np.uint8)
    img[hs // 2:, :, 1] = (np.arange(ws) * 255 // max(ws - 1, 1)).astype(np.uint8)[None, :]   # a ramp as well
    return img
