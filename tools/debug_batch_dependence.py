"""Debug tool: does frame 0 of a net's output depend on the batch it is run in? Compares against the FP32 oracle too."""
import sys
import numpy as np, torch
from openpose_b200 import synth
from openpose_b200.net import NetB200
from oracle import ref_net
model = sys.argv[1] if len(sys.argv) > 1 else "hand"
h = int(sys.argv[2]) if len(sys.argv) > 2 else 96
w = int(sys.argv[3]) if len(sys.argv) > 3 else 96
cuda = torch.device("cuda:0")
proto, cm, ref = ref_net.build(model, 4321)
x1 = synth.frames_to_net_input(synth.frames(1, h, w, 7))
want = ref.forward(x1, keep=["net_output"])["net_output"][0]
net = NetB200(proto, cm); net.initializationOnThread()
outs = {}
for n in (1, 2, 3, 4, 5, 8):
    x = np.zeros((n, 3, h, w), np.float32); x[0] = x1[0]
    if n > 1:
        x[1:] = synth.frames_to_net_input(synth.frames(n - 1, h, w, 9))
    net.forwardPass(torch.from_numpy(x).to(cuda))
    o = net.getOutputBlobArray().cpu().numpy()[0].copy()
    outs[n] = o
    print("batch %d: max-abs vs oracle %.4g, vs batch-1 output %.4g, rms %.4g" % (n, np.abs(o - want).max(), np.abs(o - outs[1]).max(), float(np.sqrt((want ** 2).mean()))))
net.close()
