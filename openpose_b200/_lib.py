"""ctypes loader for the C-ABI product library (openpose_b200/lib/libopb200.so, header include/opb200.h).

There is NO fallback: if the library is missing the import of any op fails loudly with instructions.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# OPB200_LIB_PATH: A/B runs of two builds of the SAME product library on one box (tools/ab_libs.sh); never a fallback
LIB_PATH = os.environ.get("OPB200_LIB_PATH") or os.path.join(_HERE, "lib", "libopb200.so")

OK = 0
SEMANTICS_CPU, SEMANTICS_CUDA = 0, 1
POSE_BODY_25, POSE_COCO_18 = 0, 1


class Opb200Error(RuntimeError):
    pass


class ConnectParams(C.Structure):
    _fields_ = [("pose_model", C.c_int), ("max_peaks", C.c_int),
                ("inter_min_above_threshold", C.c_float), ("inter_threshold", C.c_float),
                ("min_subset_cnt", C.c_int), ("min_subset_score", C.c_float),
                ("default_nms_threshold", C.c_float), ("scale_factor", C.c_float),
                ("maximize_positives", C.c_int), ("semantics", C.c_int)]


_I4 = C.c_int * 4
_lib = None

# every symbol include/opb200.h declares: (name, restype, argtypes)
SYMBOLS = {
    "opb200_last_error": (C.c_char_p, []),
    "opb200_version": (C.c_int, []),
    "opb200_device_count": (C.c_int, []),
    "opb200_resize_merge": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(_I4), C.POINTER(C.c_float),
                                      C.c_int, _I4, C.c_void_p]),
    "opb200_nms_scratch_bytes": (C.c_size_t, [_I4, _I4]),
    "opb200_nms": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float, _I4, _I4,
                             C.c_int, C.c_void_p]),
    "opb200_nms_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float, _I4, _I4, C.c_int,
                                    C.c_void_p]),
    "opb200_maximum_scratch_bytes": (C.c_size_t, [_I4, _I4]),
    "opb200_maximum": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, _I4, _I4, C.c_void_p]),
    "opb200_preprocess": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                    C.c_float, C.c_void_p]),
    "opb200_crop_warp": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                   C.c_void_p]),
    "opb200_heatmaps_export_channels": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int]),
    "opb200_heatmaps_export": (C.c_int, [C.c_void_p, C.c_void_p, _I4, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                         C.c_void_p]),
    "opb200_keep_top_people": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "opb200_keypoint_scale_and_offset": (C.c_int, [C.POINTER(C.c_float), C.c_int, C.c_double, C.c_double, C.c_int,
                                                   C.c_int]),
    "opb200_scale_keypoints": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float,
                                         C.c_float, C.c_float, C.c_void_p]),
    "opb200_hand_face_rectangles": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                              C.c_int, C.c_void_p]),
    "opb200_connect_scratch_bytes": (C.c_size_t, [C.POINTER(ConnectParams), C.c_int]),
    "opb200_connect": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                 C.c_void_p, _I4, C.POINTER(ConnectParams), C.c_void_p]),
    "opb200_paf_scores": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, _I4, C.POINTER(ConnectParams),
                                    C.c_void_p]),
}


def register(symbols: dict) -> None:
    """Other modules (net) add their symbols before the first lib() call or bind lazily after."""
    SYMBOLS.update(symbols)
    if _lib is not None:
        _bind(_lib, symbols)


def _bind(lib, symbols) -> None:
    for name, (res, args) in symbols.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export it
        fn.restype = res
        fn.argtypes = args


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise Opb200Error(
                "libopb200.so not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(or `make -C openpose_b200/csrc`). There is no CPU fallback on this path.")
        handle = C.CDLL(LIB_PATH)
        _bind(handle, SYMBOLS)
        _lib = handle
    return _lib


def check(status: int) -> None:
    if status != OK:
        raise Opb200Error(lib().opb200_last_error().decode("utf-8", "replace"))


def i4(shape) -> "_I4":
    s = list(shape)
    assert len(s) == 4, s
    return _I4(*[int(v) for v in s])
