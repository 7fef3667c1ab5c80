"""Debug tool: time of the first net layers via a truncated BODY_25 graph (conv1_1 only as net output)."""
import sys, os, tempfile
import torch
from openpose_b200 import net as netmod, synth, modelgen, caffe_io
B = int(sys.argv[1]) if len(sys.argv) > 1 else 44
text = ('input: "image"\ninput_dim: 1\ninput_dim: 3\ninput_dim: 16\ninput_dim: 16\n'
        'layer { name: "conv1_1" type: "Convolution" bottom: "image" top: "conv1_1" convolution_param { num_output: 64 pad: 1 kernel_size: 3 } }\n'
        'layer { name: "relu1_1" type: "ReLU" bottom: "conv1_1" top: "conv1_1" }\n'
        'layer { name: "c2" type: "Convolution" bottom: "conv1_1" top: "net_output" convolution_param { num_output: 8 pad: 0 kernel_size: 1 } }\n')
d = tempfile.mkdtemp()
nd = caffe_io.load_netdef(text)
blobs = modelgen.random_init(nd, 1)
open(os.path.join(d, "n.prototxt"), "w").write(text)
modelgen.save_caffemodel(os.path.join(d, "n.caffemodel"), nd, blobs)
dev = torch.device("cuda:0")
n = netmod.NetB200(os.path.join(d, "n.prototxt"), os.path.join(d, "n.caffemodel"))
n.initializationOnThread()
x = torch.from_numpy(synth.frames_to_net_input(synth.frames(B, 368, 656, 100))).to(dev)
for _ in range(3):
    n.forwardPass(x)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(10):
    n.forwardPass(x)
b.record()
torch.cuda.synchronize()
print("conv1_1 + 1x1(64->8) at batch %d: %.3f ms per forward" % (B, a.elapsed_time(b) / 10))
