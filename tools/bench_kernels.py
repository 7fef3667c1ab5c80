"""Debug tool: device-time of single post-processing kernels at given shapes (CUDA events, L2-cold inputs)."""
import sys
import torch
from openpose_b200 import post
from openpose_b200.op_input import preprocess_gpu


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


def main():
    dev = torch.device("cuda:0")
    for name, n, c in (("hand", 64, 22), ("face", 32, 71)):
        src = torch.rand((n, c, 368, 368), device=dev)
        dst = torch.empty((n, 1, c - 1, 3), device=dev)
        ms = timeit(lambda: post.maximum_gpu(dst, src))
        gb = n * (c - 1) * 368 * 368 * 4 / 1e9
        print("maximum %s x%d: %.3f ms  %.0f GB/s" % (name, n, ms, gb / ms * 1e3))
        small = torch.rand((n, c, 46, 46), device=dev)
        ms = timeit(lambda: post.resize_and_merge_gpu(src, [small]))
        gb = n * c * (368 * 368 + 46 * 46) * 4 / 1e9
        print("resize x8 %s x%d: %.3f ms  %.0f GB/s" % (name, n, ms, gb / ms * 1e3))
    for hs, ws, n in ((1080, 1920, 44), (720, 1280, 44), (480, 640, 44), (368, 656, 44)):
        frames = torch.randint(0, 256, (n, hs, ws, 3), dtype=torch.uint8, device=dev)
        dst = torch.empty((n, 3, 368, 656), device=dev)
        scale = min(655.0 / (ws - 1), 367.0 / (hs - 1))
        ms = timeit(lambda: preprocess_gpu(dst, frames, scale))
        gb = n * (hs * ws * 3 + 3 * 368 * 656 * 4) / 1e9
        print("preprocess %dx%d x%d: %.3f ms  %.0f GB/s" % (ws, hs, n, ms, gb / ms * 1e3))
    # BASELINE.json config 2: BODY_25 multi-scale (scale_number 4, scale_gap 0.25): four net outputs -> one heat-map blob,
    # then nms + connector on planted maps with P people
    import numpy as np
    import bench as B
    sizes = [(46, 82), (34, 60), (22, 40), (10, 20)]
    srcs = [torch.rand((1, 78, h, w), device=dev) for h, w in sizes]
    big = torch.empty((16, 78, 368, 656), device=dev)
    ms = timeit(lambda: [post.resize_and_merge_gpu(big[i:i + 1], srcs) for i in range(16)]) / 16
    gb = 4 * (78 * 368 * 656 + sum(78 * h * w for h, w in sizes)) / 1e9
    print("resizeAndMerge 4 scales -> 78x368x656, per frame: %.4f ms  %.0f GB/s" % (ms, gb / ms * 1e3))
    for people in (1, 8, 32):
        outs = torch.from_numpy(B.planted_net_outputs(16, people, 5)).to(dev)
        pp = B.PostPath(16, dev)
        post.resize_and_merge_gpu(pp.heat, [outs])
        ms_n = timeit(lambda: post.nms_gpu(pp.peaks, pp.heat, 0.05, (B.NMS_OFFSET, B.NMS_OFFSET)))
        ms_c = timeit(lambda: post.connect_body_parts_gpu(pp.heat, pp.peaks, pp.params, capacity=256,
                                                          out=(pp.kp, pp.sc, pp.cnt)))
        print("planted %2d people x16 frames: nms %.3f ms %.0f GB/s, connector %.3f ms (%d people found in frame 0)" % (
            people, ms_n, 16 * B.NMS_BYTES / 1e6 / ms_n, ms_c, int(pp.cnt[0].item())))
    heat = torch.rand((16, 78, 368, 656), device=dev)
    for mode, name in ((8, "NoScale"), (7, "UnsignedChar"), (3, "ZeroToOne")):
        ms = timeit(lambda: post.heatmaps_export_gpu(heat, 25, True, 52, 7, mode, to_host=False))
        gb = 2 * heat.numel() * 4 / 1e9
        print("heatmaps_export %s x16 -> device: %.3f ms  %.0f GB/s" % (name, ms, gb / ms * 1e3))
    one = heat[:1]
    ms = timeit(lambda: post.heatmaps_export_gpu(one, 25, True, 52, 7, 3, to_host=True), reps=10)
    print("heatmaps_export 1 frame -> pinned host: %.3f ms  %.1f GB/s over the link" % (ms, one.numel() * 4 / 1e6 / ms))
    ms = timeit(lambda: one.clamp(0, 1).cpu(), reps=10)
    print("torch clamp + .cpu() 1 frame (pageable): %.3f ms" % ms)


if __name__ == "__main__":
    main()
