"""Experiment: one net at batch B vs two nets at batch B/2 on two streams (do the per-layer bubbles overlap?)."""
import sys
import torch
from openpose_b200 import net as netmod, synth

B = int(sys.argv[1]) if len(sys.argv) > 1 else 44
dev = torch.device("cuda:0")
x = torch.from_numpy(synth.frames_to_net_input(synth.frames(B, 368, 656, 100))).to(dev)


def run(nets, xs, streams, reps=6):
    def once():
        for n, xi, s in zip(nets, xs, streams):
            with torch.cuda.stream(s):
                n.forward_device(xi)
    for _ in range(3):
        once()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        once()
    for s in streams:
        torch.cuda.current_stream().wait_stream(s)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


one = netmod.bench_net(B, 368, 656, dev)
ms1 = run([one], [x], [torch.cuda.Stream()])
print("1 x %d: %.3f ms  %.0f frames/s" % (B, ms1, B / ms1 * 1e3))
h = B // 2
two = [netmod.bench_net(h, 368, 656, dev), netmod.bench_net(B - h, 368, 656, dev)]
ms2 = run(two, [x[:h].contiguous(), x[h:].contiguous()], [torch.cuda.Stream(), torch.cuda.Stream()])
print("2 x %d: %.3f ms  %.0f frames/s" % (h, ms2, B / ms2 * 1e3))
