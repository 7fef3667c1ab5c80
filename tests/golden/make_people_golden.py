"""Generates tests/golden/people_golden.npz from the REFERENCE's own object code (oracle/_ref: handDetector.cpp,
faceDetector.cpp, keepTopNPeople.cpp, keypointScaler.cpp, keypoint.cpp compiled unmodified by oracle/Makefile).
Run in the build container:  python tests/golden/make_people_golden.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.post import Ref  # noqa: E402
from tests.cases_people import PEOPLE_CASES, SCALE_CASES, TOP_N, make_people  # noqa: E402


def main():
    ref = Ref()
    out = {}
    for name, model, parts, people, seed in PEOPLE_CASES:
        kp, sc = make_people(parts, people, seed)
        out[name + "/hands"] = ref.detect_hands(kp, model)
        out[name + "/faces"] = ref.detect_faces(kp, model)
        for n in TOP_N:
            out["%s/top%d" % (name, n)] = ref.keep_top_people(kp, sc, n)
        for i, (mode, s_io, s_no, prod) in enumerate(SCALE_CASES):
            out["%s/scale%d" % (name, i)] = ref.scale_keypoints(kp, mode, s_io, s_no, prod)
    path = os.path.join(ROOT, "tests", "golden", "people_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
