"""Debug tool: write-only / copy bandwidth ceilings of this box next to resize8Kernel (all on 44 BODY_25 frames)."""
import torch
from openpose_b200 import post

dev = torch.device("cuda:0")


def timeit(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


n = 44
heat = torch.empty((n, 78, 368, 656), device=dev)
heat2 = torch.empty_like(heat)
small = torch.rand((n, 78, 46, 82), device=dev)
gb = heat.numel() * 4 / 1e9
ms = timeit(lambda: heat.fill_(1.0))
print("fill_ %.2f GB: %.3f ms  %.0f GB/s written" % (gb, ms, gb / ms * 1e3))
ms = timeit(lambda: heat.zero_())
print("zero_ (memset) %.2f GB: %.3f ms  %.0f GB/s written" % (gb, ms, gb / ms * 1e3))
ms = timeit(lambda: heat2.copy_(heat))
print("copy_ %.2f GB: %.3f ms  %.0f GB/s read+written" % (gb, ms, 2 * gb / ms * 1e3))
ms = timeit(lambda: post.resize_and_merge_gpu(heat, [small]))
print("resize x8 x%d: %.3f ms  %.0f GB/s (algorithmic)" % (n, ms, (gb + small.numel() * 4 / 1e9) / ms * 1e3))
peaks = torch.zeros((n, 25, 128, 3), device=dev)
ms = timeit(lambda: post.nms_gpu(peaks, heat, 0.05, (0.25, 0.25)))
print("nms x%d on upsampled uniform noise: %.3f ms  %.0f GB/s" % (n, ms, n * 25 * 368 * 656 * 4 / 1e9 / ms * 1e3))
s = heat[:, :25].sum()
ms = timeit(lambda: heat[:, :25].sum())
print("torch sum over the 25 part planes (read-only): %.3f ms  %.0f GB/s" % (ms, n * 25 * 368 * 656 * 4 / 1e9 / ms * 1e3))
