"""Generates tests/golden/crop_warp.npz: outputs of cv2.warpAffine (OpenCV -- the reference's dependency for the
hand / face crop, handExtractorCaffe.cpp:44-76) for seeded frames and the matrices listed in tests/cases_crop.py.
Run in the build container:  python tests/golden/make_crop_golden.py   (cv2 4.13.0 at the time of writing)
"""
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests.cases_crop import CROP_CASES, crop_frame_input  # noqa: E402


def main():
    out = {"cv2_version": np.array([int(v) for v in cv2.__version__.split(".")[:3]], np.int32)}
    for name, seed, (hs, ws), size, m in CROP_CASES:
        img = crop_frame_input(seed, hs, ws)
        out[name] = cv2.warpAffine(img, np.asarray(m, np.float64), size, flags=cv2.INTER_LINEAR | cv2.WARP_INVERSE_MAP,
                                   borderMode=cv2.BORDER_CONSTANT, borderValue=(0, 0, 0))
    path = os.path.join(ROOT, "tests", "golden", "crop_warp.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
