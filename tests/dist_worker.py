"""Worker of tests/test_dist_cpu.py: one rank of a world_size-2 gloo job exercising bench.py's multi-rank
plumbing on CPU (rank env as torchrun sets it). Prints one JSON line."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    import bench
    from openpose_b200 import modelgen, synth
    os.environ["OPB200_DIST_BACKEND"] = "gloo"
    rank, world, local = bench.dist_init(int(os.environ["WORLD_SIZE"]))
    dev = torch.device("cpu")
    bench.barrier(world)
    # both ranks race to create the same model files in a fresh directory (the bench does this on a fresh box)
    proto, cm, nd, blobs = modelgen.build_files("hand", 4321, os.environ["OPB200_TEST_WORKDIR"])
    bench.barrier(world)
    size = os.path.getsize(cm)
    # frame sharding: every rank draws its own frames (weak scaling), nothing is exchanged
    frames = synth.frames(2, 32, 48, 100 + rank)
    digest = float(frames.astype("float64").sum())
    # whole-job rate = all units / slowest rank
    units, seconds = 10.0 * (rank + 1), 0.5 * (rank + 1)
    rate = bench.whole_job_rate(units, seconds, world, dev)
    out = {"rank": rank, "world": world, "size": size, "digest": digest, "rate": rate,
           "max": bench.max_over_ranks(float(rank), world, dev), "sum": bench.sum_over_ranks(float(rank + 1), world, dev)}
    print("RESULT " + json.dumps(out), flush=True)
    import torch.distributed as dist
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
