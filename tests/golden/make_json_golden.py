"""Writes tests/golden/json_golden.npz: the bytes the reference's OWN object code (oracle/_ref: fileStream.cpp,
jsonOfstream.cpp, cocoJsonSaver.cpp, peopleJsonSaver.cpp compiled unmodified) writes for the seeded cases of
tests/cases_json.py. Run in the build container (needs /root/reference): python tests/golden/make_json_golden.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.post import Ref  # noqa: E402
from tests import cases_json  # noqa: E402


def main():
    ref = Ref()
    out = {}
    for name, (kv, cand) in cases_json.people_cases().items():
        for hr in (0, 1):
            out["people/%s/%d" % (name, hr)] = np.frombuffer(ref.save_people_json(kv, cand, bool(hr)), np.uint8)
    for name, c in cases_json.coco_cases().items():
        for hr in (0, 1):
            files = ref.coco_json(c["frames"], c["names"], c["parts"], c["pose_model"], bool(hr), c["variants"], c["fmt"])
            for fn, data in files.items():
                out["coco/%s/%d/%s" % (name, hr, fn)] = np.frombuffer(data, np.uint8)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "json_golden.npz"), **out)
    print("wrote %d entries" % len(out))


if __name__ == "__main__":
    main()
