"""Supplementary measurement: net-forward throughput of the four OpenPose graphs (BASELINE.json configs 1 and 3) with
seeded random-init weights. usage: PYTHONPATH=. python tools/bench_models.py [steps]"""
import json
import sys
import torch
from openpose_b200 import net as netmod, synth

STEPS = int(sys.argv[1]) if len(sys.argv) > 1 else 10
dev = torch.device("cuda:0")
CASES = [("body_25", 44, 368, 656), ("coco", 64, 368, 656), ("hand", 64, 368, 368), ("face", 64, 368, 368)]
for model, n, h, w in CASES:
    x = torch.from_numpy(synth.frames_to_net_input(synth.frames(n, h, w, 7))).to(dev)
    net = netmod.bench_net(n, h, w, dev, model=model)
    for _ in range(3):
        net.forward_device(x)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(STEPS):
        net.forward_device(x)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / STEPS
    info = net.net.info()
    print(json.dumps({"model": model, "batch": n, "net_input": [h, w], "ms_per_forward": round(ms, 3),
                      "frames_or_crops_per_s": round(n / ms * 1e3, 1),
                      "gflop_per_unit": round(info.flops_per_forward / n * 1e-9, 2),
                      "conv_tflops": round(info.flops_per_forward / ms * 1e-9, 1),
                      "launches": info.kernel_launches_per_forward}))
    del net
    torch.cuda.empty_cache()
