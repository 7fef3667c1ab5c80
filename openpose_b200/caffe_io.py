"""Pure-Python Caffe prototxt / caffemodel I/O (no protobuf runtime, so it also works on the GPU box).

Host-side tooling, not the inference path: the product loads models with its own C++ reader
(csrc/net/caffe_format.cpp), which the tests check *against* this one. Used to write the seeded random-init
caffemodels of bench.py / smoke() (openpose_b200/modelgen.py) and, through oracle/caffe_io.py, by the oracle.

Formats (reference = /root/reference, files relative to it):
  * prototxt text format: the subset of protobuf text format used by
    models/pose/body_25/pose_deploy.prototxt (scalars, strings, enums, nested
    messages, '#' comments, repeated fields).
  * caffemodel binary: protobuf wire format of NetParameter, field numbers from
    3rdparty/caffe/src/caffe/proto/caffe.proto:
      NetParameter   name=1 input=3 input_dim=4 layer=100           (caffe.proto:65-100)
      LayerParameter name=1 type=2 bottom=3 top=4 blobs=7            (caffe.proto:327-408)
      BlobProto      num=1 channels=2 height=3 width=4 data=5 shape=7 (caffe.proto:6-22)
      BlobShape      dim=1 (packed int64)
  Weight matching is by layer *name* as in caffe::Net::CopyTrainedLayersFrom
  (3rdparty/caffe/src/caffe/net.cpp).
"""
from __future__ import annotations

import re
import struct
from dataclasses import dataclass, field
from typing import Dict, List, Tuple

import numpy as np

# --------------------------------------------------------------------------------------
# prototxt (text format) parser
# --------------------------------------------------------------------------------------

_TOKEN = re.compile(r'\s*(?:(#[^\n]*)|([A-Za-z_][A-Za-z0-9_]*)|("(?:[^"\\]|\\.)*")|'
                    r"('(?:[^'\\]|\\.)*')|([-+]?[0-9.][0-9a-zA-Z.+-]*)|([{}:<>\[\],;]))")


def _tokenize(text: str) -> List[str]:
    out, pos, n = [], 0, len(text)
    while pos < n:
        m = _TOKEN.match(text, pos)
        if not m:
            if text[pos:].strip() == "":
                break
            raise ValueError("prototxt: cannot tokenize at %r" % text[pos:pos + 30])
        pos = m.end()
        if m.group(1) is not None:      # comment
            continue
        out.append(m.group(0).strip())
    return out


def _parse_message(tokens: List[str], i: int, closer: str | None) -> Tuple[Dict[str, list], int]:
    """Returns ({field: [values...]}, next index). Values are str / float / int / dict."""
    msg: Dict[str, list] = {}
    while i < len(tokens):
        t = tokens[i]
        if closer is not None and t == closer:
            return msg, i + 1
        if t in (",", ";"):
            i += 1
            continue
        name = t
        i += 1
        if i < len(tokens) and tokens[i] == ":":
            i += 1
        t = tokens[i]
        if t in ("{", "<"):
            sub, i = _parse_message(tokens, i + 1, "}" if t == "{" else ">")
            msg.setdefault(name, []).append(sub)
        else:
            msg.setdefault(name, []).append(_scalar(t))
            i += 1
    if closer is not None:
        raise ValueError("prototxt: missing closing %s" % closer)
    return msg, i


def _scalar(t: str):
    if t[0] in "\"'":
        return bytes(t[1:-1], "utf-8").decode("unicode_escape")
    if re.fullmatch(r"[-+]?\d+", t):
        return int(t)
    try:
        return float(t)
    except ValueError:
        return t  # enum identifier / true / false


def parse_prototxt(text: str) -> Dict[str, list]:
    msg, _ = _parse_message(_tokenize(text), 0, None)
    return msg


@dataclass
class Layer:
    name: str
    type: str
    bottoms: List[str]
    tops: List[str]
    params: Dict[str, object] = field(default_factory=dict)


@dataclass
class NetDef:
    name: str
    inputs: List[str]
    input_dim: List[int]
    layers: List[Layer]


def _first(d, key, default=None):
    v = d.get(key)
    return v[0] if v else default


def load_netdef(path_or_text: str) -> NetDef:
    text = path_or_text
    if "\n" not in path_or_text and not path_or_text.lstrip().startswith("name"):
        with open(path_or_text, "r") as f:
            text = f.read()
    msg = parse_prototxt(text)
    layers = []
    for L in msg.get("layer", []):
        p: Dict[str, object] = {}
        t = _first(L, "type")
        if t == "Convolution":
            cp = _first(L, "convolution_param", {})
            p = dict(num_output=_first(cp, "num_output"), kernel_size=_first(cp, "kernel_size", 1),
                     pad=_first(cp, "pad", 0), stride=_first(cp, "stride", 1),
                     group=_first(cp, "group", 1), dilation=_first(cp, "dilation", 1),
                     bias_term=_first(cp, "bias_term", "true") not in ("false", False, 0))
        elif t == "Pooling":
            pp = _first(L, "pooling_param", {})
            p = dict(pool=_first(pp, "pool", "MAX"), kernel_size=_first(pp, "kernel_size"),
                     stride=_first(pp, "stride", 1), pad=_first(pp, "pad", 0))
        elif t == "Concat":
            cp = _first(L, "concat_param", {})
            p = dict(axis=_first(cp, "axis", 1))
        elif t == "PReLU":
            pp = _first(L, "prelu_param", {})
            p = dict(channel_shared=_first(pp, "channel_shared", "false") in ("true", True, 1))
        layers.append(Layer(_first(L, "name"), t, list(L.get("bottom", [])), list(L.get("top", [])), p))
    return NetDef(_first(msg, "name", ""), list(msg.get("input", [])),
                  [int(x) for x in msg.get("input_dim", [])], layers)


# --------------------------------------------------------------------------------------
# protobuf wire format helpers
# --------------------------------------------------------------------------------------

def _varint(v: int) -> bytes:
    out = bytearray()
    v &= (1 << 64) - 1
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _key(field_no: int, wire: int) -> bytes:
    return _varint((field_no << 3) | wire)


def _ld(field_no: int, payload: bytes) -> bytes:
    return _key(field_no, 2) + _varint(len(payload)) + payload


def _read_varint(buf: memoryview, i: int) -> Tuple[int, int]:
    shift, v = 0, 0
    while True:
        b = buf[i]
        i += 1
        v |= (b & 0x7F) << shift
        if not b & 0x80:
            return v, i
        shift += 7


def _iter_fields(buf: memoryview):
    i, n = 0, len(buf)
    while i < n:
        k, i = _read_varint(buf, i)
        fno, wire = k >> 3, k & 7
        if wire == 0:
            v, i = _read_varint(buf, i)
            yield fno, wire, v
        elif wire == 1:
            yield fno, wire, buf[i:i + 8]
            i += 8
        elif wire == 2:
            ln, i = _read_varint(buf, i)
            yield fno, wire, buf[i:i + ln]
            i += ln
        elif wire == 5:
            yield fno, wire, buf[i:i + 4]
            i += 4
        else:
            raise ValueError("caffemodel: unsupported wire type %d" % wire)


def encode_blob(arr: np.ndarray) -> bytes:
    arr = np.ascontiguousarray(arr, dtype="<f4")
    shape = b"".join(_varint(int(d)) for d in arr.shape)
    return _ld(7, _ld(1, shape)) + _ld(5, arr.tobytes())


def decode_blob(buf: memoryview) -> np.ndarray:
    legacy = {}
    shape = None
    chunks: List[np.ndarray] = []
    for fno, wire, v in _iter_fields(buf):
        if fno == 7 and wire == 2:
            dims = []
            for f2, w2, v2 in _iter_fields(v):
                if f2 == 1 and w2 == 2:
                    j = 0
                    while j < len(v2):
                        d, j = _read_varint(v2, j)
                        dims.append(d)
                elif f2 == 1 and w2 == 0:
                    dims.append(v2)
            shape = dims
        elif fno == 5 and wire == 2:
            chunks.append(np.frombuffer(v, dtype="<f4"))
        elif fno == 5 and wire == 5:
            chunks.append(np.frombuffer(v, dtype="<f4"))
        elif fno in (1, 2, 3, 4) and wire == 0:
            legacy[fno] = v
    data = np.concatenate(chunks) if chunks else np.zeros(0, "<f4")
    if shape is None:
        shape = [legacy.get(k, 1) for k in (1, 2, 3, 4)] if legacy else [data.size]
    return data.reshape(shape).astype(np.float32)


def write_caffemodel(path: str, net_name: str, layer_blobs: List[Tuple[str, str, List[np.ndarray]]]) -> None:
    """layer_blobs: [(layer name, layer type, [blob arrays])] in network order."""
    out = bytearray(_ld(1, net_name.encode()))
    for name, ltype, blobs in layer_blobs:
        body = _ld(1, name.encode()) + _ld(2, ltype.encode())
        for b in blobs:
            body += _ld(7, encode_blob(b))
        out += _ld(100, body)
    with open(path, "wb") as f:
        f.write(bytes(out))


def read_caffemodel(path: str) -> Dict[str, List[np.ndarray]]:
    with open(path, "rb") as f:
        buf = memoryview(f.read())
    out: Dict[str, List[np.ndarray]] = {}
    for fno, wire, v in _iter_fields(buf):
        if fno == 100 and wire == 2:
            name, blobs = None, []
            for f2, w2, v2 in _iter_fields(v):
                if f2 == 1 and w2 == 2:
                    name = bytes(v2).decode()
                elif f2 == 7 and w2 == 2:
                    blobs.append(decode_blob(v2))
            if name is not None and blobs:
                out[name] = blobs
    return out
