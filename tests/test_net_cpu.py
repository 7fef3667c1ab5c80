"""CPU tests of the net path's host logic (no GPU): generated graphs vs the reference's prototxt files,
the oracle's layer semantics (KATs restated from Caffe's own tests), the C++ prototxt/caffemodel readers and
graph planner (through the host-only C-ABI entry opb200_net_describe) vs the Python oracle readers."""
import hashlib
import json
import os

import numpy as np
import pytest
import torch

from openpose_b200 import net as netmod
from openpose_b200 import modelgen, protogen
from openpose_b200._lib import Opb200Error
from oracle import caffe_io, ref_net

REF = "/root/reference"
SIG_FILE = os.path.join(os.path.dirname(__file__), "golden", "prototxt_signatures.json")


def _signature(nd: caffe_io.NetDef) -> str:
    rows = [[L.name, L.type, L.bottoms, L.tops, {k: (v if not isinstance(v, bool) else int(v))
                                                  for k, v in sorted(L.params.items())}] for L in nd.layers]
    blob = json.dumps([nd.name, nd.inputs, rows], sort_keys=True).encode()
    return hashlib.sha256(blob).hexdigest()


@pytest.mark.parametrize("model", protogen.MODELS)
def test_generated_prototxt_equals_reference(model):
    gen = caffe_io.load_netdef(protogen.generate(model))
    with open(SIG_FILE) as f:
        sigs = json.load(f)
    assert _signature(gen) == sigs[model]            # committed signature of the reference's file
    path = os.path.join(REF, protogen.REFERENCE_FILES[model])
    if os.path.exists(path):                         # and the file itself, where the reference tree exists
        ref = caffe_io.load_netdef(path)
        assert _signature(ref) == sigs[model]
        assert [(L.name, L.type, L.bottoms, L.tops, L.params) for L in ref.layers] == \
               [(L.name, L.type, L.bottoms, L.tops, L.params) for L in gen.layers]


# ---- oracle KATs (specifications restated from 3rdparty/caffe/src/caffe/test/) ---------------------------
def _one_layer_net(layer_text, cin):
    return caffe_io.load_netdef('input: "image"\ninput_dim: 1\ninput_dim: %d\ninput_dim: 4\ninput_dim: 4\n' % cin + layer_text)


def test_oracle_conv_matches_naive_definition():
    """test_convolution_layer.cpp:22-147 (caffe_conv reference loop): cross-correlation with zero padding."""
    rng = np.random.default_rng(0)
    nd = _one_layer_net('layer { name: "c" type: "Convolution" bottom: "image" top: "net_output" '
                        'convolution_param { num_output: 4 pad: 1 kernel_size: 3 } }', 3)
    w = rng.standard_normal((4, 3, 3, 3)).astype(np.float32)
    b = rng.standard_normal(4).astype(np.float32)
    x = rng.standard_normal((2, 3, 5, 6)).astype(np.float32)
    got = ref_net.RefNet(nd, {"c": [w, b]}).forward(x)["net_output"]
    want = np.zeros((2, 4, 5, 6), np.float64)
    xp = np.pad(x, ((0, 0), (0, 0), (1, 1), (1, 1)))
    for n in range(2):
        for o in range(4):
            for y in range(5):
                for xx in range(6):
                    want[n, o, y, xx] = (xp[n, :, y:y + 3, xx:xx + 3].astype(np.float64) * w[o]).sum() + b[o]
    assert np.abs(got - want).max() < 1e-5


def test_oracle_sobel_kat():
    """test_convolution_layer.cpp:498-589: a 3x3 Sobel filter equals its separable (3x1)*(1x3) form."""
    x = np.random.default_rng(1).standard_normal((1, 1, 8, 9)).astype(np.float32)
    sobel = np.array([[-1, 0, 1], [-2, 0, 2], [-1, 0, 1]], np.float32).reshape(1, 1, 3, 3)
    nd = _one_layer_net('layer { name: "c" type: "Convolution" bottom: "image" top: "net_output" '
                        'convolution_param { num_output: 1 pad: 1 kernel_size: 3 } }', 1)
    got = ref_net.RefNet(nd, {"c": [sobel, np.zeros(1, np.float32)]}).forward(x)["net_output"][0, 0]
    xp = np.pad(x[0, 0], 1)
    col = xp[:-2] + 2 * xp[1:-1] + xp[2:]                   # (1,2,1)^T along y
    want = col[:, 2:] - col[:, :-2]                         # (-1,0,1) along x
    assert np.abs(got - want).max() < 1e-5


def test_oracle_prelu_closed_form_and_ceil_pool():
    """test_neuron_layer.cpp:124-150,786-803 (PReLU = max(x,0)+a_c*min(x,0)); test_pooling_layer.cpp (2x2 stride 2
    MAX pooling uses the CEIL output size and clips windows at the border)."""
    x = np.random.default_rng(2).standard_normal((1, 3, 5, 7)).astype(np.float32)
    nd = _one_layer_net('layer { name: "p" type: "PReLU" bottom: "image" top: "image" }\n'
                        'layer { name: "q" type: "Pooling" bottom: "image" top: "net_output" '
                        'pooling_param { pool: MAX kernel_size: 2 stride: 2 } }', 3)
    a = np.array([0.25, 0.1, 0.5], np.float32)
    got = ref_net.RefNet(nd, {"p": [a]}).forward(x)["net_output"]
    y = np.maximum(x, 0) + a[None, :, None, None] * np.minimum(x, 0)
    assert got.shape == (1, 3, 3, 4)                        # ceil(5/2), ceil(7/2)
    for i in range(3):
        for j in range(4):
            assert np.array_equal(got[0, :, i, j], y[0, :, 2 * i:2 * i + 2, 2 * j:2 * j + 2].max(axis=(1, 2)))


# ---- C++ readers + planner (host-only ABI entry) vs the oracle readers -----------------------------------
@pytest.mark.parametrize("model", protogen.MODELS)
def test_cpp_planner_covers_every_layer(model):
    proto, cm, ref = ref_net.build(model, 1234)
    text = netmod.describe(proto, cm)
    lines = text.strip().splitlines()
    convs = [l.split() for l in lines if l.startswith("conv ")]
    nd = ref.netdef
    want = [L for L in nd.layers if L.type == "Convolution"]
    assert [c[1] for c in convs] == [L.name for L in want]
    blobs = caffe_io.read_caffemodel(cm)
    for c, L in zip(convs, want):
        kv = dict(t.split("=", 1) for t in c[2:])
        assert int(kv["cout"]) == L.params["num_output"] == blobs[L.name][0].shape[0]
        assert int(kv["cin"]) == blobs[L.name][0].shape[1]
        segs = [tuple(int(v) for v in s.split(":")) for s in kv["segs"].strip("[]").split(",") if s]
        if L.bottoms[0] != "image":
            assert sum(s[2] for s in segs) == int(kv["cin"])          # concat segments cover the input channels
        assert all(s[1] % 8 == 0 for s in segs)                      # 16-byte aligned TMA bases
    # three 2x2 poolings: stand-alone launches or folded into the epilogue of the conv that feeds them
    fused = sum(int(dict(t.split("=", 1) for t in c[2:])["pool"]) for c in convs)
    assert sum(1 for l in lines if l.startswith("pool ")) + fused == 3
    assert lines[-1].startswith("output net_output C=%d " % protogen.OUTPUT_CHANNELS[model])
    # every member of net_output is written in fp32 by a conv epilogue (no convert pass needed)
    assert not any(l.startswith("convert ") for l in lines)
    assert sum(int(dict(t.split("=", 1) for t in c[2:])["out32"]) >= 0 for c in convs) == (2 if model in ("body_25", "coco") else 1)


def test_cpp_reader_error_paths(tmp_path):
    proto, cm, _ = ref_net.build("hand", 1234)
    with pytest.raises(Opb200Error, match="output blob is a nullptr"):
        netmod.describe(proto, cm, "nope")
    bad = tmp_path / "bad.prototxt"
    bad.write_text('input: "image"\ninput_dim: 1\ninput_dim: 3\ninput_dim: 8\ninput_dim: 8\n'
                   'layer { name: "x" type: "Deconvolution" bottom: "image" top: "net_output" }\n')
    with pytest.raises(Opb200Error, match="unsupported layer type Deconvolution"):
        netmod.describe(str(bad), cm)
    trunc = tmp_path / "trunc.caffemodel"
    trunc.write_bytes(open(cm, "rb").read()[:1000])
    with pytest.raises(Opb200Error, match="caffemodel"):
        netmod.describe(proto, str(trunc))
    strided = tmp_path / "strided.prototxt"
    strided.write_text('input: "image"\ninput_dim: 1\ninput_dim: 3\ninput_dim: 8\ninput_dim: 8\n'
                       'layer { name: "conv1_1" type: "Convolution" bottom: "image" top: "net_output" '
                       'convolution_param { num_output: 64 kernel_size: 3 pad: 1 stride: 2 } }\n')
    with pytest.raises(Opb200Error, match="only stride-1"):
        netmod.describe(str(strided), cm)


def test_random_init_is_seeded_and_caffemodel_roundtrips(tmp_path):
    nd = caffe_io.load_netdef(protogen.generate("body_25"))
    a, b = ref_net.random_init(nd, 7), ref_net.random_init(nd, 7)
    assert all(np.array_equal(x, y) for k in a for x, y in zip(a[k], b[k]))
    p = str(tmp_path / "m.caffemodel")
    ref_net.save_caffemodel(p, nd, a)
    back = caffe_io.read_caffemodel(p)
    assert set(back) == set(a)
    assert all(np.array_equal(x, y) for k in a for x, y in zip(a[k], back[k]))
    assert sum(x.size for k in a for x in a[k]) > 26_000_000          # BODY_25: 26.15 M parameters


# ---- the net oracle's convolution / pooling arithmetic against the reference's bundled Caffe code ----
def _ref_caffe():
    import ctypes
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libopref_caffe.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libopref_caffe.so not built (needs /root/reference at build time)")
    L = ctypes.CDLL(path)
    fp = ctypes.POINTER(ctypes.c_float)
    L.refcaffe_conv.argtypes = [fp, fp, fp, fp] + [ctypes.c_int] * 6
    L.refcaffe_conv.restype = None
    L.refcaffe_max_pool.argtypes = [fp, fp] + [ctypes.c_int] * 6
    L.refcaffe_max_pool.restype = ctypes.c_int
    return L, fp


@pytest.mark.parametrize("cin,cout,h,w,k", [(3, 8, 9, 11, 3), (16, 12, 7, 10, 7), (24, 20, 6, 5, 1), (5, 4, 1, 1, 3),
                                            (64, 32, 12, 14, 3)])
def test_oracle_conv_equals_caffe_im2col_gemm(cin, cout, h, w, k):
    """oracle/ref_net.py's Convolution (torch conv2d, padding k//2) == Caffe's own im2col_cpu (object code of
    3rdparty/caffe/src/caffe/util/im2col.cpp) followed by the GEMM + bias of base_conv_layer.cpp:258-281, to 1e-5."""
    import torch.nn.functional as F
    L, fp = _ref_caffe()
    rng = np.random.default_rng(cin * 100 + k)
    x = rng.normal(0, 1, (cin, h, w)).astype(np.float32)
    wt = (rng.normal(0, 1, (cout, cin, k, k)) / np.sqrt(cin * k * k)).astype(np.float32)
    b = rng.uniform(-0.1, 0.1, cout).astype(np.float32)
    out = np.empty((cout, h, w), np.float32)
    L.refcaffe_conv(out.ctypes.data_as(fp), x.ctypes.data_as(fp), wt.ctypes.data_as(fp), b.ctypes.data_as(fp), cin, cout, h,
                    w, k, k // 2)
    want = F.conv2d(torch.from_numpy(x)[None], torch.from_numpy(wt), torch.from_numpy(b), padding=k // 2)[0].numpy()
    assert np.abs(out - want).max() <= 1e-5 * max(1.0, np.abs(want).max())


@pytest.mark.parametrize("h,w", [(8, 8), (9, 11), (7, 6), (1, 5), (46, 82), (23, 41)])
def test_oracle_max_pool_equals_caffe_forward(h, w):
    """oracle/ref_net.py's Pooling (max_pool2d 2x2 / 2, ceil_mode) == PoolingLayer::Forward_cpu as restated from
    pooling_layer.cpp:91-111,141-175 (ceil shape rule, windows clipped to the image), bit for bit."""
    import torch.nn.functional as F
    L, fp = _ref_caffe()
    rng = np.random.default_rng(h * 100 + w)
    x = rng.normal(0, 1, (3, h, w)).astype(np.float32)
    shape = L.refcaffe_max_pool(None, x.ctypes.data_as(fp), 3, h, w, 2, 2, 0)
    ph, pw = shape >> 16, shape & 0xFFFF
    out = np.empty((3, ph, pw), np.float32)
    L.refcaffe_max_pool(out.ctypes.data_as(fp), x.ctypes.data_as(fp), 3, h, w, 2, 2, 0)
    want = F.max_pool2d(torch.from_numpy(x)[None], 2, 2, 0, ceil_mode=True)[0].numpy()
    assert want.shape == out.shape and np.array_equal(out, want)


def test_prototxt_hostile_nesting_and_anisotropic_pooling(tmp_path):
    """the C++ reader caps message nesting (no stack overflow on a file of '{') and rejects non-square pooling windows
    instead of silently planning them as 1x1 (the conv path already rejects non-square kernels)."""
    _, cm, _, _ = __import__("openpose_b200.modelgen", fromlist=["build_files"]).build_files("body_25", 1234)
    bad = tmp_path / "nest.prototxt"
    bad.write_text("layer " + "{ a " * 20000)
    with pytest.raises(Opb200Error, match="nested deeper"):
        netmod.describe(str(bad), cm)
    pool = tmp_path / "pool.prototxt"
    pool.write_text('input: "image"\ninput_dim: 1\ninput_dim: 3\ninput_dim: 16\ninput_dim: 16\n'
                    'layer { name: "p" type: "Pooling" bottom: "image" top: "net_output" '
                    'pooling_param { pool: MAX kernel_h: 2 kernel_w: 3 stride: 2 } }\n')
    with pytest.raises(Opb200Error, match="non-square pooling"):
        netmod.describe(str(pool), cm)


@pytest.mark.parametrize("model,shape,level_hw,flat,sep", [
    ("body_25", (47, 368, 656), (46, 82), 1, 1),      # 3x3 convs only: one separator column
    ("coco", (64, 368, 656), (46, 82), 1, 3),         # 7x7 convs (pad 3) read the level: three
    ("hand", (32, 96, 96), (12, 12), 1, 3),
    ("face", (32, 96, 96), (12, 12), 1, 3),
    ("hand", (16, 368, 368), (46, 46), 0, 0),         # 46 + 3 columns per frame no longer save a tile against 3 x 16
    ("hand", (2, 96, 96), (12, 12), 0, 0),            # too few frames for the interleaved layout to save tiles
    ("body_25", (1, 368, 656), (46, 82), 0, 0),
])
def test_interleaved_levels_have_separators_as_wide_as_the_padding(model, shape, level_hw, flat, sep):
    """Frame-interleaved ("flat") levels rely on zero separator columns between frames as the horizontal padding, so a level
    read by a conv with pad p needs p of them (regression: one column under the 7x7 layers of COCO / hand / face made their
    filters read the neighbouring frame). Host-only, through opb200_net_describe_shape."""
    proto, cm, _, _ = modelgen.build_files(model, 1234)
    levels = netmod.describe_shape(proto, cm, *shape)
    assert levels[0]["h"] == shape[1] and levels[0]["w"] == shape[2] and levels[0]["flat"] == 0
    hits = [lv for lv in levels if (lv["h"], lv["w"]) == level_hw]
    assert len(hits) == 1 and hits[0]["convs"] > 0
    assert hits[0]["flat"] == flat and hits[0]["sep"] == sep
    for lv in levels:
        if lv["flat"]:
            assert lv["sep"] >= max(1, lv["max_pad"])
