"""One process, N GPUs: frames/s of opb200_engine_* with batches dealt round-robin over the devices and results returned
in frame order (the in-library counterpart of the reference's one-extractor-per-GPU topology).
usage: PYTHONPATH=. python tools/engine_multi_gpu.py [n_gpus] [steps] [batch]"""
import json
import sys
import time

import torch

from openpose_b200 import modelgen, synth
from openpose_b200.engine import PoseEngine


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else torch.cuda.device_count()
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 60
    batch = int(sys.argv[3]) if len(sys.argv) > 3 else 44
    proto, cm, _, _ = modelgen.build_files("body_25", 1234)
    eng = PoseEngine(proto, cm, gpus=list(range(n)), batch=batch, lanes=2, capacity=256)
    frames = synth.frames(batch, 720, 1280, 300)
    fid = 0
    for _ in range(2 * n):                       # fill every lane of every device
        eng.acquire()[:] = frames
        eng.commit(fid, batch)
        fid += batch
    expect = 0
    for _ in range(3 * n):                       # warm-up in steady state
        first = eng.next(copy=False)[0]
        assert first == expect, (first, expect)
        expect += batch
        eng.acquire()
        eng.commit(fid, batch)
        fid += batch
    for d in range(n):
        torch.cuda.synchronize(d)
    t0 = time.perf_counter()
    for _ in range(steps * n):
        first = eng.next(copy=False)[0]
        assert first == expect, (first, expect)  # frame order across devices
        expect += batch
        eng.acquire()
        eng.commit(fid, batch)
        fid += batch
    wall = time.perf_counter() - t0
    while eng.in_flight():
        first = eng.next(copy=False)[0]
        assert first == expect
        expect += batch
    eng.close()
    print(json.dumps({"tool": "engine_multi_gpu", "n_gpus": n, "batch": batch, "steps_per_gpu": steps,
                      "frames_per_s": steps * n * batch / wall, "ms_per_batch_per_gpu": wall / steps * 1e3,
                      "mode": "one process, one worker thread per device, u8 1280x720 frames from pinned staging, frame order checked"}))


if __name__ == "__main__":
    main()
