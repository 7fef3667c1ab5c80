"""TEST INFRASTRUCTURE (oracle) -- Caffe prototxt / caffemodel I/O for the oracle net.

The pure-Python reader/writer lives in openpose_b200/caffe_io.py (host tooling that also writes the bench's
random-init caffemodels); the oracle uses the same file formats, so it re-exports it. The product's inference
path never touches either: it has its own C++ reader (csrc/net/caffe_format.cpp), checked against this one.
"""
from openpose_b200.caffe_io import *  # noqa: F401,F403
from openpose_b200.caffe_io import (Layer, NetDef, decode_blob, encode_blob, load_netdef, parse_prototxt,  # noqa: F401
                                    read_caffemodel, write_caffemodel)
