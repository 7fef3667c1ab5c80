"""include/opb200.h is strict C99 and libopb200.so links from plain C: tests/c/abi_example.c compiles with
`gcc -std=c99 -pedantic -Werror` (CPU), and on a GPU its output equals the Python mirror's on the same frames."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "openpose_b200", "lib")
LOOSE = dict(nms_threshold=0.02, inter_threshold=0.0, inter_min_above=0.3, min_subset_cnt=2, min_subset_score=0.0)


def _build(tmp_path):
    exe = str(tmp_path / "abi_example")
    cmd = ["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "c", "abi_example.c"), "-o", exe, "-L", LIBDIR, "-lopb200",
           "-Wl,-rpath," + LIBDIR]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_c_example_compiles_and_links_as_c99(tmp_path):
    from openpose_b200 import _lib
    _lib.lib()                                     # the library is built
    exe = _build(tmp_path)
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 2 and "usage" in r.stderr          # no GPU call without arguments


@pytest.mark.gpu
def test_c_example_equals_python_engine(tmp_path, cuda):
    from openpose_b200 import modelgen, synth
    from openpose_b200.engine import PoseEngine
    proto, cm, _, _ = modelgen.build_files("body_25", 1234)
    w, h, net, batch, n = 320, 180, (160, 96), 3, 8
    frames = synth.frames(n, h, w, 77)
    raw = tmp_path / "frames.raw"
    frames.tofile(str(raw))
    eng = PoseEngine(proto, cm, gpus=[0], net_size=net, frame_size=(w, h), batch=batch, lanes=2, capacity=64, overrides=LOOSE)
    want = []
    for i in range(0, n, batch):
        eng.submit(i, frames[i:i + batch])
        first, kp, sc, cnt = eng.next()
        for f in range(cnt.shape[0]):
            line = "frame %d people %d" % (first + f, int(cnt[f]))
            if cnt[f] > 0:
                line += " score %.9g kp %s %s %s" % (sc[f, 0], float(kp[f, 0, 0, 0]).hex(), float(kp[f, 0, 0, 1]).hex(),
                                                    float(kp[f, 0, 0, 2]).hex())
            want.append(line)
    scale_in, scale_out = eng.cfg.scale_input_to_net, eng.scale_net_to_output
    eng.close()
    exe = _build(tmp_path)
    r = subprocess.run([exe, proto, cm, str(raw), str(n), str(w), str(h), str(net[0]), str(net[1]), repr(float(scale_in)),
                        repr(float(scale_out)), str(batch)], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    got = r.stdout.strip().splitlines()
    assert len(got) == n and sum(int(l.split()[3]) for l in got) > 0

    def parse(line):
        t = line.split()
        vals = [float.fromhex(v) for v in t[7:10]] if len(t) > 4 else []
        return t[1], t[3], (float(t[5]) if len(t) > 4 else None), vals
    for g, wl in zip(got, want):
        pg, pw = parse(g), parse(wl)
        assert pg[0] == pw[0] and pg[1] == pw[1] and pg[3] == pw[3], (g, wl)
        if pw[2] is not None:
            assert np.float32(pg[2]) == np.float32(pw[2]), (g, wl)
