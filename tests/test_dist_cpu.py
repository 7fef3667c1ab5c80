"""world_size-2 gloo test (CPU) of the N>1 path's host logic: rank env handling, barrier / max / sum helpers,
whole-job rate (sum of units / slowest rank), per-rank frame shards, and two ranks racing to write the same
seeded model files (what happens on a fresh box when torchrun starts bench.py on every GPU)."""
import json
import os
import socket
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_two_rank_gloo_plumbing(tmp_path):
    port = _free_port()
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), OPB200_TEST_WORKDIR=str(tmp_path / "models"), OMP_NUM_THREADS="2")
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "dist_worker.py")], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    res = []
    for p in procs:
        out, err = p.communicate(timeout=240)
        assert p.returncode == 0, err[-2000:]
        res.append(json.loads([l for l in out.splitlines() if l.startswith("RESULT ")][-1][7:]))
    res.sort(key=lambda r: r["rank"])
    assert [r["rank"] for r in res] == [0, 1] and all(r["world"] == 2 for r in res)
    assert res[0]["size"] == res[1]["size"] > 1_000_000            # one intact caffemodel, seen by both
    assert res[0]["digest"] != res[1]["digest"]                    # different frames per rank
    assert all(r["max"] == 1.0 and r["sum"] == 3.0 for r in res)
    assert all(abs(r["rate"] - (10.0 + 20.0) / 1.0) < 1e-9 for r in res)   # 30 units / slowest rank's 1.0 s
    assert not [f for f in os.listdir(tmp_path / "models") if ".tmp" in f]
