"""Debug tool: device time of the x8 single-scale resizeAndMerge at the headline shape (44 x 78 x 46 x 82 -> 368 x 656),
L2-cold (the destination, 3.3 GB, is far larger than L2). OPB200_RESIZE8_STREAM=0 selects the shared-memory kernel."""
import sys
import torch
from openpose_b200 import post


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 44
    dev = torch.device("cuda:0")
    for c, hs, ws in ((78, 46, 82), (22, 46, 46), (71, 46, 46)):
        src = torch.rand((n, c, hs, ws), device=dev)
        dst = torch.empty((n, c, hs * 8, ws * 8), device=dev)
        for _ in range(3):
            post.resize_and_merge_gpu(dst, [src])
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(20):
            post.resize_and_merge_gpu(dst, [src])
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 20
        gb = (dst.numel() + src.numel()) * 4 / 1e9
        print("resize8 %dx%dx%dx%d: %.3f ms  %.0f GB/s" % (n, c, hs, ws, ms, gb / ms * 1e3))


if __name__ == "__main__":
    main()
