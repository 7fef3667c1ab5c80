"""Debug tool: nms on the dense maps a random-weight net emits (~127 peaks per channel) vs planted 8-person maps.
usage: bench_nms_dense.py [batch] [once]   ('once': one launch per case after warm-up, for ncu)"""
import sys
import numpy as np
import torch
import bench as B
from openpose_b200 import net as netmod, post, synth


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    once = len(sys.argv) > 2
    dev = torch.device("cuda:0")
    x = torch.from_numpy(synth.frames_to_net_input(synth.frames(n, 368, 656, 100))).to(dev)
    net = netmod.bench_net(n, 368, 656, dev)
    dense = net.forward_device(x).clone()
    planted = torch.from_numpy(B.planted_net_outputs(n, 8, 5)).to(dev)
    pp = B.PostPath(n, dev)
    for name, outs in (("planted", planted), ("dense", dense)):
        post.resize_and_merge_gpu(pp.heat, [outs])
        for _ in range(3):
            post.nms_gpu(pp.peaks, pp.heat, 0.05, (B.NMS_OFFSET, B.NMS_OFFSET))
        torch.cuda.synchronize()
        reps = 1 if once else 20
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            post.nms_gpu(pp.peaks, pp.heat, 0.05, (B.NMS_OFFSET, B.NMS_OFFSET))
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / reps
        cnt = pp.peaks[:, :, 0, 0]
        print("nms %-8s x%d: %.3f ms  %.0f GB/s   peaks per channel: mean %.1f max %d" % (
            name, n, ms, n * B.NMS_BYTES / 1e6 / ms, float(cnt.mean()), int(cnt.max())))


if __name__ == "__main__":
    main()
