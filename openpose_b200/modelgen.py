"""Seeded random-init model files for bench.py / smoke() / the tests (there is no network for the trained
caffemodels): the four architectures of protogen.py with weights that keep activations O(1) end to end.
Host tooling only; the files it writes are loaded by the C++ reader like any trained caffemodel."""
import os
from typing import Dict, List, Optional, Tuple

import numpy as np

from . import caffe_io, protogen


def random_init(netdef: caffe_io.NetDef, seed: int, out_rms: float = 0.15) -> Dict[str, List[np.ndarray]]:
    """He-normal convs (gain for the activation that follows), small biases, PReLU slopes U(0.1, 0.3); convs
    with no activation after them (the stage outputs) are scaled so their outputs have RMS ~ out_rms,
    i.e. heat-map-like magnitudes (the 2e-2 max-abs tolerance of the north star is stated for such maps)."""
    rng = np.random.default_rng(seed)
    chan = {netdef.inputs[0]: netdef.input_dim[1] if len(netdef.input_dim) > 1 else 3}
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


def build_files(model: str, seed: int = 1234, workdir: Optional[str] = None
                ) -> Tuple[str, str, caffe_io.NetDef, Dict[str, List[np.ndarray]]]:
    """Writes <workdir>/<model>_seed<seed>.prototxt + .caffemodel once (safe when several ranks race for it:
    each writes its own temporary file and renames it into place) and returns (proto, caffemodel, netdef, blobs)."""
    workdir = workdir or os.path.join(os.environ.get("TMPDIR", "/tmp"), "opb200_models")
    os.makedirs(workdir, exist_ok=True)
    text = protogen.generate(model)
    proto = os.path.join(workdir, "%s_seed%d.prototxt" % (model, seed))
    cm = os.path.join(workdir, "%s_seed%d.caffemodel" % (model, seed))
    nd = caffe_io.load_netdef(text)
    blobs = random_init(nd, seed)
    if not (os.path.exists(proto) and os.path.exists(cm)):
        tmp = ".tmp%d" % os.getpid()
        with open(proto + tmp, "w") as f:
            f.write(text)
        os.replace(proto + tmp, proto)
        save_caffemodel(cm + tmp, nd, blobs)
        os.replace(cm + tmp, cm)
    return proto, cm, nd, blobs
