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


# The seeded random-init generator and the caffemodel writer are host tooling shared with bench.py / smoke():
# they live in openpose_b200/modelgen.py (the product never imports the oracle, the oracle may use the tooling).
from openpose_b200.modelgen import random_init, save_caffemodel  # noqa: E402,F401


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
    from openpose_b200 import modelgen
    proto, cm, nd, blobs = modelgen.build_files(model, seed, workdir)
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
