"""Debug tool: device time of the 4-scale resizeAndMerge of BASELINE config 2 (one frame per call, as the reference's
multi-scale flow issues it). OPB200_RESIZE_MERGE_STREAM=0 selects the band kernel."""
import torch
from openpose_b200 import post


def main():
    dev = torch.device("cuda:0")
    sizes = [(46, 82), (34, 60), (22, 40), (10, 20)]
    ratios = [1.0, 0.75, 0.5, 0.25]
    for ns in (4, 2):
        srcs = [torch.rand((1, 78, h, w), device=dev) for h, w in sizes[:ns]]
        big = torch.empty((16, 78, 368, 656), device=dev)
        def run():
            for i in range(16):
                post.resize_and_merge_gpu(big[i:i + 1], srcs, ratios[:ns])
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(10):
            run()
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 160
        gb = 4 * (78 * 368 * 656 + sum(78 * h * w for h, w in sizes[:ns])) / 1e9
        print("resizeAndMerge %d scales -> 78x368x656, per frame: %.4f ms  %.0f GB/s" % (ns, ms, gb / ms * 1e3))


if __name__ == "__main__":
    main()
