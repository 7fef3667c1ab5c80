"""Generates tests/golden/post_golden.npz from the REFERENCE's own object code (oracle/_ref, built from
/root/reference sources by oracle/Makefile). Run in the build container:  python tests/golden/make_golden.py
Inputs are re-generated from seeds by tests/cases.py; a CRC of every input is stored to catch generator drift.
"""
import os
import sys
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.post import Ref, model_dims  # noqa: E402
from tests.cases import POST_CASES, case_id, make_case  # noqa: E402


def main():
    ref = Ref()
    out = {}
    for case in POST_CASES:
        model = case[0]
        parts = model_dims(model)[0]
        maps, _ = make_case(case)
        cid = case_id(case)
        peaks = ref.nms(maps, 0.05, parts, off=(0.25, 0.25))
        # zero the stale rows so the fixture is deterministic
        for c in range(parts):
            peaks[c, int(peaks[c, 0, 0]) + 1:] = 0
            peaks[c, 0, 1:] = 0
        out[cid + "/crc"] = np.array([zlib.crc32(maps.tobytes())], np.uint32)
        out[cid + "/peaks"] = peaks
        for mp in (0, 1):
            kp, sc = ref.connect(maps, peaks, model, scale=1.959, maximize_positives=bool(mp))
            out["%s/kp%d" % (cid, mp)] = kp
            out["%s/sc%d" % (cid, mp)] = sc
    path = os.path.join(ROOT, "tests", "golden", "post_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()


def make_prototxt_signatures():
    """tests/golden/prototxt_signatures.json: sha256 of the canonical layer list of the REFERENCE's four deploy
    prototxts (needs /root/reference); tests/test_net_cpu.py compares openpose_b200.protogen's output to them."""
    import json
    from oracle import caffe_io
    from openpose_b200 import protogen
    from tests.test_net_cpu import SIG_FILE, _signature
    sigs = {m: _signature(caffe_io.load_netdef("/root/reference/" + protogen.REFERENCE_FILES[m]))
            for m in protogen.MODELS}
    with open(SIG_FILE, "w") as f:
        json.dump(sigs, f, indent=1)
