"""Experiment: post-processing of step i (HBM bound) on a second stream while the net of step i+1 (tensor bound)
runs. usage: PYTHONPATH=. python tools/bench_pipeline.py [batch] [steps] [net_priority] [post_priority]"""
import sys
import torch
import bench
from openpose_b200 import net as netmod, synth

B = int(sys.argv[1]) if len(sys.argv) > 1 else 44
STEPS = int(sys.argv[2]) if len(sys.argv) > 2 else 12
PN = int(sys.argv[3]) if len(sys.argv) > 3 else 0
PP = int(sys.argv[4]) if len(sys.argv) > 4 else 0
dev = torch.device("cuda:0")
x = torch.from_numpy(synth.frames_to_net_input(synth.frames(B, 368, 656, 100))).to(dev)
net = netmod.bench_net(B, 368, 656, dev)
post = bench.PostPath(B, dev)


def serial(n):
    for _ in range(n):
        post.run(net.forward_device(x))


s_net = torch.cuda.Stream(device=dev, priority=PN)
s_post = torch.cuda.Stream(device=dev, priority=PP)
stash = None


def pipelined(n):
    global stash
    copied = torch.cuda.Event()
    for i in range(n):
        with torch.cuda.stream(s_net):
            if i:
                s_net.wait_event(copied)            # the previous output has been moved out of the net's blob
            out = net.forward_device(x)
            done = torch.cuda.Event()
            done.record()
        with torch.cuda.stream(s_post):
            s_post.wait_event(done)
            if stash is None:
                stash = torch.empty_like(out)
            stash.copy_(out)
            copied = torch.cuda.Event()
            copied.record()
            post.run(stash)
    torch.cuda.current_stream().wait_stream(s_post)
    torch.cuda.current_stream().wait_stream(s_net)


def timed(fn, n):
    fn(3)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    fn(n)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / n
    return ms


ms = timed(serial, STEPS)
print("serial   : %.3f ms/step  %.0f frames/s" % (ms, B / ms * 1e3))
ms = timed(pipelined, STEPS)
print("pipelined: %.3f ms/step  %.0f frames/s (priorities net %d post %d)" % (ms, B / ms * 1e3, PN, PP))
kp, sc, cnt = post.kp.clone(), post.sc.clone(), post.cnt.clone()
serial(1)
torch.cuda.synchronize()
n = cnt.cpu()
same = bool((cnt == post.cnt).all()) and all(
    torch.equal(kp[i, :n[i]], post.kp[i, :n[i]]) and torch.equal(sc[i, :n[i]], post.sc[i, :n[i]]) for i in range(B))
print("pipelined result == serial result:", same)
