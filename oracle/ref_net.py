"""TEST INFRASTRUCTURE (oracle) -- CPU FP32 restatement of the Caffe inference subset the OpenPose nets use.

The reference's forward pass is caffe::Net::ForwardFrom(0) (src/openpose/net/netCaffe.cpp:248) over the
layers of the deploy prototxt. Caffe itself cannot be built here (boost/glog/protobuf/BLAS absent, see
SURVEY.md 8(c)), so this file restates the layer semantics with torch CPU ops, FP32 throughout:
  Convolution  3rdparty/caffe/src/caffe/layers/conv_layer.cpp:8-40, base_conv_layer.cpp:258-281
               -> F.conv2d(x, W (Cout,Cin,kh,kw), b, stride, padding)          (cross-correlation, like Caffe)
  ReLU         relu_layer.cpp                       -> F.relu (negative_slope 0)
  PReLU        prelu_layer.cpp:67-89                -> max(x,0) + a_c * min(x,0), per channel
  Pooling MAX  pooling_layer.cpp:91-97,141-175      -> F.max_pool2d(k, s, ceil_mode=True) (pad 0: ceil size, clipped)
  Concat       concat_layer.cpp (axis 1)            -> torch.cat(dim=1)
Weights are matched by layer name (net.cpp CopyTrainedLayersFrom); blobs[0] = W, blobs[1] = bias; PReLU
blobs[0] = slopes. Parity note: "parity unpinned by the reference's own tests" -- the reference holds no golden
vectors for its nets (no trained weights in-tree); this restatement is checked against Caffe's documented
layer definitions via the KATs in tests/test_ref_net.py (restated from 3rdparty/caffe/src/caffe/test/).

Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference) may import this file.
"""
from __future__ import annotations

import os
import time
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch
import torch.nn.functional as F

from . import caffe_io


def random_init(netdef: caffe_io.NetDef, seed: int, out_rms: float = 0.15) -> Dict[str, List[np.ndarray]]:
    """Seeded random-init weights that keep activations O(1) through the whole graph:
    He-normal convs (gain for the activation that follows), small biases, PReLU slopes U(0.1, 0.3); convs
    with no activation after them (the stage outputs) are scaled so their outputs have RMS ~ out_rms,
    i.e. heat-map-like magnitudes (the 2e-2 max-abs tolerance of the north star is stated for such maps)."""
    rng = np.random.default_rng(seed)
    chan = {netdef.inputs[0]: netdef.input_dim[1] if len(netdef.input_dim) > 1 else 3}
    follows: Dict[str, str] = {}
    for L in netdef.layers:
        if L.type in ("ReLU", "PReLU") and L.bottoms == L.tops:
            follows.setdefault(L.bottoms[0] + "@" + str(id(L)), L.type)
    blobs: Dict[str, List[np.ndarray]] = {}
    layers = netdef.layers
    for i, L in enumerate(layers):
        if L.type == "Convolution":
            cin, cout, k = chan[L.bottoms[0]], int(L.params["num_output"]), int(L.params["kernel_size"])
            act = None
            for A in layers[i + 1:]:
                if A.bottoms == [L.tops[0]] and A.tops == [L.tops[0]] and A.type in ("ReLU", "PReLU"):
                    act = A.type
                    break
                if L.tops[0] in A.bottoms:
                    break
            fan_in = cin * k * k
            if act == "ReLU":
                std = np.sqrt(2.0 / fan_in)
            elif act == "PReLU":
                std = np.sqrt(2.0 / (1.0 + 0.2 ** 2) / fan_in)
            else:
                std = out_rms * np.sqrt(1.0 / fan_in)
            w = (rng.standard_normal((cout, cin, k, k)) * std).astype(np.float32)
            b = rng.uniform(-0.05, 0.05, size=(cout,)).astype(np.float32) * (1.0 if act else out_rms)
            blobs[L.name] = [w, b]
            chan[L.tops[0]] = cout
        elif L.type == "PReLU":
            blobs[L.name] = [rng.uniform(0.1, 0.3, size=(chan[L.bottoms[0]],)).astype(np.float32)]
        elif L.type == "Concat":
            chan[L.tops[0]] = sum(chan[b] for b in L.bottoms)
        else:
            chan[L.tops[0]] = chan[L.bottoms[0]]
    return blobs


def save_caffemodel(path: str, netdef: caffe_io.NetDef, blobs: Dict[str, List[np.ndarray]]) -> None:
    recs = [(L.name, L.type, blobs[L.name]) for L in netdef.layers if L.name in blobs]
    caffe_io.write_caffemodel(path, netdef.name or "net", recs)


class RefNet:
    """Executes a NetDef with FP32 torch CPU ops. forward() returns every blob by top name."""

    def __init__(self, netdef: caffe_io.NetDef, blobs: Dict[str, List[np.ndarray]]):
        self.netdef = netdef
        self.w = {k: [torch.from_numpy(np.ascontiguousarray(a)) for a in v] for k, v in blobs.items()}

    @torch.no_grad()
    def forward(self, x: np.ndarray, keep: Optional[List[str]] = None, bf16_activations: bool = False
                ) -> Dict[str, np.ndarray]:
        """bf16_activations=True additionally rounds every conv input/weight to bf16 (fp32 accumulate), i.e. the
        arithmetic the B200 path uses -- for separating rounding effects from bugs in the tests."""
        t: Dict[str, torch.Tensor] = {self.netdef.inputs[0]: torch.from_numpy(np.ascontiguousarray(x, np.float32))}
        rb = (lambda a: a.to(torch.bfloat16).to(torch.float32)) if bf16_activations else (lambda a: a)
        for L in self.netdef.layers:
            if L.type == "Convolution":
                p = L.params
                w, b = self.w[L.name][0], (self.w[L.name][1] if len(self.w[L.name]) > 1 else None)
                y = F.conv2d(rb(t[L.bottoms[0]]), rb(w), b, stride=int(p["stride"]), padding=int(p["pad"]),
                             dilation=int(p["dilation"]), groups=int(p["group"]))
            elif L.type == "ReLU":
                y = F.relu(t[L.bottoms[0]])
            elif L.type == "PReLU":
                a = self.w[L.name][0]
                y = F.prelu(t[L.bottoms[0]], a if a.numel() > 1 else a.reshape(1))
            elif L.type == "Pooling":
                p = L.params
                assert p["pool"] in ("MAX", 0), "only MAX pooling occurs in the OpenPose nets"
                y = F.max_pool2d(t[L.bottoms[0]], int(p["kernel_size"]), int(p["stride"]), int(p["pad"]),
                                 ceil_mode=True)
            elif L.type == "Concat":
                y = torch.cat([t[b] for b in L.bottoms], dim=int(L.params.get("axis", 1)))
            elif L.type == "Input":
                continue
            else:
                raise NotImplementedError("layer type %s" % L.type)
            t[L.tops[0]] = y
        names = keep if keep is not None else list(t.keys())
        return {k: t[k].numpy() for k in names}


def build(model: str, seed: int = 1234, workdir: Optional[str] = None) -> Tuple[str, str, RefNet]:
    """Writes <workdir>/<model>.prototxt + .caffemodel (seeded random init) and returns (proto, model, RefNet)."""
    from openpose_b200 import protogen
    workdir = workdir or os.path.join(os.environ.get("TMPDIR", "/tmp"), "opb200_models")
    os.makedirs(workdir, exist_ok=True)
    text = protogen.generate(model)
    proto = os.path.join(workdir, "%s_seed%d.prototxt" % (model, seed))
    cm = os.path.join(workdir, "%s_seed%d.caffemodel" % (model, seed))
    nd = caffe_io.load_netdef(text)
    blobs = random_init(nd, seed)
    if not (os.path.exists(proto) and os.path.exists(cm)):
        with open(proto, "w") as f:
            f.write(text)
        save_caffemodel(cm + ".tmp", nd, blobs)
        os.replace(cm + ".tmp", cm)
    return proto, cm, RefNet(nd, blobs)


_cache: Dict[Tuple[str, int], RefNet] = {}


def timed_forward_body25(x: np.ndarray, threads: int, seed: int = 1234):
    """bench.py's CPU leg for the net: (net_output, seconds) for the batch x on `threads` host threads."""
    torch.set_num_threads(max(1, threads))
    key = ("body_25", seed)
    if key not in _cache:
        _cache[key] = build("body_25", seed)[2]
    net = _cache[key]
    t0 = time.perf_counter()
    out = net.forward(x, keep=["net_output"])["net_output"]
    return out, time.perf_counter() - t0
