"""Throughput of the in-library hand / face crop engine (opb200_crop_engine_*) on the GPUs given: u8 1280x720 frames with
`crops` person crops each (BASELINE config 4: per-person crops sharded across GPUs), host buffers in, maxima out.
usage: crop_engine_bench.py [hand|face] [crops per frame] [frames] [gpu ids, comma separated]"""
import sys
import time
import numpy as np
from openpose_b200 import modelgen, synth
from openpose_b200.crop_extractors import crop_matrix
from openpose_b200.engine import CropEngine


def main():
    model = sys.argv[1] if len(sys.argv) > 1 else "hand"
    crops = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    frames = int(sys.argv[3]) if len(sys.argv) > 3 else 60
    gpus = [int(g) for g in sys.argv[4].split(",")] if len(sys.argv) > 4 else [0]
    channels = 22 if model == "hand" else 71
    proto, cm, _, _ = modelgen.build_files(model, 4321)
    eng = CropEngine(proto, cm, channels, gpus=gpus, frame_size=(1280, 720), net_side=368, batch=crops, lanes=2)
    rng = np.random.default_rng(0)
    frame = synth.frames(1, 720, 1280, 3)[0]
    mats = np.stack([crop_matrix((float(rng.uniform(0, 1000)), float(rng.uniform(0, 500)), float(rng.uniform(120, 260)),) * 1 +
                                 (0.0,), 368, mirror=bool(k % 2)) for k in range(crops)])
    for k in range(crops):   # square rectangles: (x, y, side, side)
        mats[k] = crop_matrix((float(rng.uniform(0, 1000)), float(rng.uniform(0, 500)), 200.0, 200.0), 368, bool(k % 2))
    cap = 2 * len(gpus)
    done = 0

    def pump(n):
        nonlocal done
        for i in range(n):
            if eng.in_flight() == cap:
                eng.next()
                done += 1
            eng.submit(i, frame, mats)
    pump(3 * cap)                                   # warm-up (graph capture, clocks)
    while eng.in_flight():
        eng.next()
    done = 0
    t0 = time.perf_counter()
    pump(frames)
    while eng.in_flight():
        eng.next()
        done += 1
    dt = time.perf_counter() - t0
    print('{"tool": "crop_engine_bench", "model": "%s", "gpus": %d, "crops_per_frame": %d, "frames": %d, "crops_per_s": %.1f, '
          '"frames_per_s": %.1f, "ms_per_batch": %.3f}' % (model, len(gpus), crops, frames, frames * crops / dt, frames / dt,
                                                          1e3 * dt / frames * len(gpus)))
    eng.close()


if __name__ == "__main__":
    main()
