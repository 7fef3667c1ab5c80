"""TEST INFRASTRUCTURE (oracle) -- ctypes front-ends for the two CPU checkers.

  Port  -> oracle/_build/libopb_oracle.so  (post_oracle.c, our plain-C restatement)
  Ref   -> oracle/_ref/libopref_post.so    (the reference's own object code + ref_shim.cpp)

Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference) may import this.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import List, Optional, Sequence, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(_HERE, "_build", "libopb_oracle.so")
REF_SO = os.path.join(_HERE, "_ref", "libopref_post.so")

BODY_25, COCO_18 = 0, 1
MODE_CPU, MODE_CUDA = 0, 1
MAX_PEAKS = 127  # POSE_MAX_PEOPLE, include/openpose/pose/poseParameters.hpp:14

_f32p = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int)


def build(ref: bool = True) -> None:
    """Compile the checkers (oracle/Makefile). The _ref target is a no-op without /root/reference."""
    subprocess.check_call(["make", "-s", "-C", _HERE, "oracle"] + (["ref"] if ref else []))


def _fp(a: np.ndarray):
    assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_f32p)


def _ip(a: Optional[np.ndarray]):
    if a is None:
        return None
    assert a.dtype == np.int32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_i32p)


def model_dims(model: int) -> Tuple[int, int]:
    """(number of body parts, number of limbs/PAF pairs)."""
    return {BODY_25: (25, 26), COCO_18: (18, 19)}[model]


def pose_defaults(model: int, maximize_positives: bool = False) -> dict:
    """src/openpose/pose/poseParameters.cpp:677-756 (restated; checked against Ref in the tests)."""
    mp = maximize_positives
    return dict(nms_threshold=0.02 if mp else 0.05, inter_min_above=0.75 if mp else 0.95,
                inter_threshold=0.01 if mp else 0.05, min_subset_cnt=2 if mp else 3,
                min_subset_score=0.05 if mp else 0.4)


class Port:
    """Our C restatement (post_oracle.c)."""

    def __init__(self):
        if not os.path.exists(PORT_SO):
            build(ref=False)
        L = self.lib = C.CDLL(PORT_SO)
        L.opo_nms.restype = C.c_int
        L.opo_nms.argtypes = [_f32p, _f32p, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float,
                              C.c_float, C.c_int]
        L.opo_resize_merge.restype = None
        L.opo_resize_merge.argtypes = [_f32p, C.POINTER(_f32p), _i32p, _i32p, _f32p, _f32p, C.c_int, C.c_int,
                                       C.c_int, C.c_int, C.c_int]
        L.opo_maximum.restype = None
        L.opo_maximum.argtypes = [_f32p, _f32p, C.c_int, C.c_int, C.c_int]
        L.opo_preprocess.restype = C.c_int
        L.opo_preprocess.argtypes = [_f32p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int]
        L.opo_crop_warp.restype = None
        L.opo_crop_warp.argtypes = [_f32p, C.c_void_p, C.POINTER(C.c_double), C.c_int, C.c_int, C.c_int, C.c_int]
        L.opo_heatmaps_export.restype = C.c_int
        L.opo_heatmaps_export.argtypes = [_f32p, _f32p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
        L.opo_pair_scores.restype = None
        L.opo_pair_scores.argtypes = [_f32p, _f32p, _f32p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float,
                                      C.c_float, C.c_float, C.c_int]
        L.opo_connect_cpu.restype = C.c_int
        L.opo_connect_cpu.argtypes = [_f32p, _f32p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float,
                                      C.c_int, C.c_float, C.c_float, C.c_float, C.c_int, _f32p, _f32p, _i32p,
                                      C.c_int]
        L.opo_connect_from_pair_scores.restype = C.c_int
        L.opo_connect_from_pair_scores.argtypes = [_f32p, _f32p, C.c_int, C.c_int, C.c_int, C.c_float,
                                                   C.c_float, C.c_int, _f32p, _f32p, _i32p, C.c_int]

    def nms(self, heat: np.ndarray, thr: float, channels: int, off=(0.5, 0.5), max_peaks=MAX_PEAKS,
            mode=MODE_CPU, peaks: Optional[np.ndarray] = None) -> np.ndarray:
        """heat (C>=channels, H, W) -> peaks (channels, max_peaks+1, 3). Rows beyond count keep `peaks`' content."""
        heat = np.ascontiguousarray(heat, np.float32)
        _, h, w = heat.shape
        if peaks is None:
            peaks = np.zeros((channels, max_peaks + 1, 3), np.float32)
        rc = self.lib.opo_nms(_fp(peaks), _fp(heat), thr, channels, h, w, max_peaks, off[0], off[1], mode)
        if rc:
            raise ValueError("threshold value invalid.")
        return peaks

    def resize_merge(self, srcs: Sequence[np.ndarray], ht: int, wt: int,
                     scale_ratios: Optional[Sequence[float]] = None, contract: bool = True) -> np.ndarray:
        """srcs[i] (C, h_i, w_i) -> (C, ht, wt); scale factors per resizeAndMergeBase.cu:379-380,475-477.
        contract: True = the FMA chain of the reference's CUDA build (pinned on the GPU against its kernels),
        False = cuda.hu as written (pinned on the CPU against cuda.hu compiled as host C++)."""
        srcs = [np.ascontiguousarray(s, np.float32) for s in srcs]
        n = len(srcs)
        c = srcs[0].shape[0]
        sw, sh = scale_factors([s.shape for s in srcs], ht, wt, scale_ratios)
        hs = np.array([s.shape[1] for s in srcs], np.int32)
        ws = np.array([s.shape[2] for s in srcs], np.int32)
        ptrs = (_f32p * n)(*[_fp(s) for s in srcs])
        dst = np.empty((c, ht, wt), np.float32)
        self.lib.opo_resize_merge(_fp(dst), ptrs, _ip(hs), _ip(ws), _fp(sw), _fp(sh), n, c, ht, wt, int(contract))
        return dst

    def maximum(self, heat: np.ndarray, parts: int) -> np.ndarray:
        heat = np.ascontiguousarray(heat, np.float32)
        _, h, w = heat.shape
        out = np.empty((parts, 3), np.float32)
        self.lib.opo_maximum(_fp(out), _fp(heat), parts, h, w)
        return out

    def preprocess(self, frame: np.ndarray, net_size, scale: float, contract: bool = True) -> np.ndarray:
        """frame (Hs,Ws,3) u8 BGR -> (3,Ht,Wt) fp32 net input; net_size = (Ht, Wt)."""
        frame = np.ascontiguousarray(frame, np.uint8)
        hs, ws, ch = frame.shape
        assert ch == 3
        out = np.empty((3, int(net_size[0]), int(net_size[1])), np.float32)
        rc = self.lib.opo_preprocess(_fp(out), frame.ctypes.data_as(C.c_void_p), hs, ws, out.shape[1],
                                     out.shape[2], float(scale), int(contract))
        assert rc == 0
        return out

    def crop_warp(self, frame: np.ndarray, matrix: np.ndarray, size) -> np.ndarray:
        """frame (Hs,Ws,3) u8, matrix 2x3 float64 inverse map, size = (Wt, Ht) as cv::Size -> (3,Ht,Wt) fp32."""
        frame = np.ascontiguousarray(frame, np.uint8)
        m = np.ascontiguousarray(matrix, np.float64).reshape(6)
        out = np.empty((3, int(size[1]), int(size[0])), np.float32)
        self.lib.opo_crop_warp(_fp(out), frame.ctypes.data_as(C.c_void_p), m.ctypes.data_as(C.POINTER(C.c_double)),
                               frame.shape[0], frame.shape[1], out.shape[1], out.shape[2])
        return out

    def heatmaps_export(self, heat: np.ndarray, parts: int, has_bkg: bool, pafs: int, mask: int, mode: int):
        """heat (C,H,W) of one frame -> (selected,H,W); None when the background is requested but absent."""
        heat = np.ascontiguousarray(heat, np.float32)
        _, h, w = heat.shape
        n = (parts if mask & 1 else 0) + (1 if (mask & 2) and has_bkg else 0) + (pafs if mask & 4 else 0)
        out = np.empty((n, h, w), np.float32)
        rc = self.lib.opo_heatmaps_export(_fp(out), _fp(heat), h * w, parts, int(has_bkg), pafs, mask, mode)
        if rc < 0:
            return None
        assert rc == n
        return out

    def pair_scores(self, heat, peaks, model, p=None, max_peaks=MAX_PEAKS, mode=MODE_CPU) -> np.ndarray:
        p = p or pose_defaults(model)
        heat = np.ascontiguousarray(heat, np.float32)
        peaks = np.ascontiguousarray(peaks, np.float32)
        _, h, w = heat.shape
        cube = np.empty((model_dims(model)[1], max_peaks, max_peaks), np.float32)
        self.lib.opo_pair_scores(_fp(cube), _fp(heat), _fp(peaks), model, w, h, max_peaks,
                                 p["inter_threshold"], p["inter_min_above"], p["nms_threshold"], mode)
        return cube

    def connect(self, heat, peaks, model, p=None, scale=1.0, maximize_positives=False, max_peaks=MAX_PEAKS,
                capacity=1024):
        """connectBodyPartsCpu -> (keypoints (P, parts, 3), scores (P,), rows (P, parts+1) int32)."""
        p = p or pose_defaults(model, maximize_positives)
        heat = np.ascontiguousarray(heat, np.float32)
        peaks = np.ascontiguousarray(peaks, np.float32)
        _, h, w = heat.shape
        parts = model_dims(model)[0]
        kp = np.zeros((capacity, parts, 3), np.float32)
        sc = np.zeros((capacity,), np.float32)
        rows = np.zeros((capacity, parts + 1), np.int32)
        n = self.lib.opo_connect_cpu(_fp(heat), _fp(peaks), model, w, h, max_peaks, p["inter_min_above"],
                                     p["inter_threshold"], int(p["min_subset_cnt"]), p["min_subset_score"],
                                     p["nms_threshold"], scale, int(maximize_positives), _fp(kp), _fp(sc),
                                     _ip(rows), capacity)
        assert 0 <= n <= capacity, n
        return kp[:n].copy(), sc[:n].copy(), rows[:n].copy()

    def connect_from_pair_scores(self, cube, peaks, model, p=None, scale=1.0, maximize_positives=False,
                                 max_peaks=MAX_PEAKS, capacity=1024):
        p = p or pose_defaults(model, maximize_positives)
        cube = np.ascontiguousarray(cube, np.float32)
        peaks = np.ascontiguousarray(peaks, np.float32)
        parts = model_dims(model)[0]
        kp = np.zeros((capacity, parts, 3), np.float32)
        sc = np.zeros((capacity,), np.float32)
        rows = np.zeros((capacity, parts + 1), np.int32)
        n = self.lib.opo_connect_from_pair_scores(_fp(cube), _fp(peaks), model, max_peaks,
                                                  int(p["min_subset_cnt"]), p["min_subset_score"], scale,
                                                  int(maximize_positives), _fp(kp), _fp(sc), _ip(rows), capacity)
        assert 0 <= n <= capacity, n
        return kp[:n].copy(), sc[:n].copy(), rows[:n].copy()


def scale_factors(shapes, ht: int, wt: int, scale_ratios=None):
    """FP32 scaleWidths/scaleHeights exactly as resizeAndMergeBase.cu:379-380,475-477 computes them."""
    n = len(shapes)
    ratios = np.asarray(scale_ratios if scale_ratios is not None else [1.0] * n, np.float32)
    w0, h0 = shapes[0][-1], shapes[0][-2]
    main_w = np.float32(wt) / np.float32(w0)
    main_h = np.float32(ht) / np.float32(h0)
    sw = np.empty(n, np.float32)
    sh = np.empty(n, np.float32)
    for i in range(n):
        r = np.float32(ratios[i] / ratios[0])
        sw[i] = main_w / r
        sh[i] = main_h / r
    return sw, sh


class Ref:
    """The reference's own object code (nmsCpu / connectBodyPartsCpu / pafPtrIntoVector...)."""

    def __init__(self):
        if not os.path.exists(REF_SO):
            if os.path.exists("/root/reference/src/openpose/net/nmsBase.cpp"):
                build(ref=True)
            if not os.path.exists(REF_SO):
                raise FileNotFoundError(REF_SO)
        L = self.lib = C.CDLL(REF_SO)
        L.ref_last_error.restype = C.c_char_p
        L.ref_nms_cpu.restype = C.c_int
        L.ref_nms_cpu.argtypes = [_f32p, _i32p, _f32p, C.c_float, _i32p, _i32p, C.c_float, C.c_float]
        L.ref_connect_cpu.restype = C.c_int
        L.ref_connect_cpu.argtypes = [_f32p, _f32p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float,
                                      C.c_int, C.c_float, C.c_float, C.c_float, C.c_int, _f32p, _f32p, C.c_int]
        L.ref_connect_from_pair_scores.restype = C.c_int
        L.ref_connect_from_pair_scores.argtypes = [_f32p, _f32p, C.c_int, C.c_int, C.c_int, C.c_float,
                                                   C.c_float, C.c_int, _f32p, _f32p, C.c_int]
        L.ref_detect_hands.restype = C.c_int
        L.ref_detect_hands.argtypes = [_f32p, C.c_int, C.c_int, C.c_int, _f32p]
        L.ref_detect_faces.restype = C.c_int
        L.ref_detect_faces.argtypes = [_f32p, C.c_int, C.c_int, C.c_int, _f32p]
        L.ref_keep_top_people.restype = C.c_int
        L.ref_keep_top_people.argtypes = [_f32p, _f32p, C.c_int, C.c_int, C.c_int, _f32p]
        L.ref_scale_keypoints.restype = C.c_int
        L.ref_scale_keypoints.argtypes = [_f32p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_int, C.c_int]
        L.ref_pose_part_pairs.argtypes = [C.c_int, C.POINTER(C.c_uint), C.c_int]
        L.ref_pose_map_index.argtypes = [C.c_int, C.POINTER(C.c_uint), C.c_int]
        L.ref_pose_defaults.argtypes = [C.c_int, C.c_int, _f32p]
        L.ref_pose_defaults.restype = None

    @staticmethod
    def available() -> bool:
        return os.path.exists(REF_SO) or os.path.exists("/root/reference/src/openpose/net/nmsBase.cpp")

    def _check(self, rc):
        if rc < 0:
            raise ValueError(self.lib.ref_last_error().decode())
        return rc

    def detect_hands(self, kp: np.ndarray, model: int) -> np.ndarray:
        """HandDetector::detectHands: kp (people, parts, 3) -> (people, 2, 4)."""
        kp = np.ascontiguousarray(kp, np.float32)
        out = np.zeros((kp.shape[0], 2, 4), np.float32)
        self._check(self.lib.ref_detect_hands(_fp(kp), kp.shape[0], kp.shape[1], model, _fp(out)))
        return out

    def detect_faces(self, kp: np.ndarray, model: int) -> np.ndarray:
        kp = np.ascontiguousarray(kp, np.float32)
        out = np.zeros((kp.shape[0], 4), np.float32)
        self._check(self.lib.ref_detect_faces(_fp(kp), kp.shape[0], kp.shape[1], model, _fp(out)))
        return out

    def keep_top_people(self, kp: np.ndarray, scores: np.ndarray, n_max: int) -> np.ndarray:
        kp = np.ascontiguousarray(kp, np.float32)
        scores = np.ascontiguousarray(scores, np.float32)
        out = np.zeros_like(kp)
        n = self._check(self.lib.ref_keep_top_people(_fp(kp), _fp(scores), kp.shape[0], kp.shape[1], n_max, _fp(out)))
        return out[:n]

    def scale_keypoints(self, kp: np.ndarray, mode: int, scale_in_out: float, scale_net_out: float, producer) -> np.ndarray:
        kp = np.array(kp, np.float32, order="C")
        self._check(self.lib.ref_scale_keypoints(_fp(kp), kp.shape[0], kp.shape[1], mode, scale_in_out, scale_net_out,
                                                 int(producer[0]), int(producer[1])))
        return kp

    def save_people_json(self, keypoint_vector, candidates=None, human_readable=True) -> bytes:
        """PeopleJsonSaver::save of the reference (own object code) into a temp directory -> the file's bytes."""
        import tempfile
        n = len(keypoint_vector)
        keep, ptrs = [], (_f32p * max(1, n))()
        dims, ndims = np.zeros((max(1, n), 3), np.int32), np.zeros((max(1, n),), np.int32)
        names = (C.c_char_p * max(1, n))()
        for i, (a, name) in enumerate(keypoint_vector):
            names[i] = name.encode()
            if a is None or np.size(a) == 0:
                continue
            a = np.ascontiguousarray(a, np.float32)
            keep.append(a)
            ptrs[i] = _fp(a)
            ndims[i] = a.ndim
            dims[i, :a.ndim] = a.shape
        n_parts = len(candidates) if candidates is not None else 0
        counts = np.zeros((max(1, n_parts),), np.int32)
        flat = np.zeros((1, 3), np.float32)
        if n_parts:
            parts = [np.asarray(c, np.float32).reshape(-1, 3) for c in candidates]
            counts[:n_parts] = [c.shape[0] for c in parts]
            flat = np.ascontiguousarray(np.concatenate(parts + [np.zeros((1, 3), np.float32)], 0), np.float32)
        self.lib.ref_save_people_json.restype = C.c_int
        self.lib.ref_save_people_json.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.POINTER(_f32p), _i32p, _i32p,
                                                  C.POINTER(C.c_char_p), _f32p, _i32p, C.c_int, C.c_int]
        with tempfile.TemporaryDirectory() as d:
            self._check(self.lib.ref_save_people_json(d.encode(), b"frame", n, ptrs, _ip(dims), _ip(ndims), names,
                                                      _fp(flat), _ip(counts), n_parts, int(human_readable)))
            with open(os.path.join(d, "frame.json"), "rb") as f:
                return f.read()

    def coco_json(self, frames, image_names, parts, pose_model=0, human_readable=True, variants=1, fmt=0,
                  variant=0) -> dict:
        """CocoJsonSaver of the reference over `frames` = [(keypoints (P,parts,3), scores (P))] -> {file suffix: bytes}."""
        import tempfile
        people = np.array([np.asarray(k).reshape(-1, parts, 3).shape[0] for k, _ in frames], np.int32)
        kp = np.ascontiguousarray(np.concatenate([np.asarray(k, np.float32).reshape(-1, parts, 3) for k, _ in frames]
                                                 + [np.zeros((1, parts, 3), np.float32)], 0))
        sc = np.ascontiguousarray(np.concatenate([np.asarray(s, np.float32).reshape(-1) for _, s in frames]
                                                 + [np.zeros((1,), np.float32)], 0))
        names = (C.c_char_p * max(1, len(frames)))()
        for i, nme in enumerate(image_names):
            names[i] = nme.encode()
        self.lib.ref_coco_json.restype = C.c_int
        self.lib.ref_coco_json.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _f32p, _f32p,
                                           _i32p, C.c_int, C.POINTER(C.c_char_p)]
        with tempfile.TemporaryDirectory() as d:
            path = os.path.join(d, "coco.json")
            self._check(self.lib.ref_coco_json(path.encode(), pose_model, int(human_readable), variants, fmt, variant,
                                               len(frames), _fp(kp), _fp(sc), _ip(people), parts, names))
            out = {}
            for fn in sorted(os.listdir(d)):
                with open(os.path.join(d, fn), "rb") as f:
                    out[fn] = f.read()
            return out

    # ---- include/openpose_private/gpu/cuda.hu compiled as host C++ (oracle/ref_cudahu_shim.cpp) ----
    def hu_resize8(self, src: np.ndarray, factor: int = 8) -> np.ndarray:
        """resize8TimesKernel's pixel walk over cuda.hu's bicubicInterpolate / bicubicInterpolate8Times."""
        src = np.ascontiguousarray(src, np.float32)
        c, h, w = src.shape
        out = np.empty((c, factor * h, factor * w), np.float32)
        self.lib.refhu_resize8.restype = None
        self.lib.refhu_resize8.argtypes = [_f32p, _f32p] + [C.c_int] * 6
        self.lib.refhu_resize8(_fp(out), _fp(src), c, h, w, factor * h, factor * w, factor)
        return out

    def hu_resize_average(self, srcs, ht: int, wt: int, scale_ratios=None) -> np.ndarray:
        """resizeAndAddAndAverageKernel's pixel walk over cuda.hu's bicubicInterpolate."""
        srcs = [np.ascontiguousarray(s, np.float32) for s in srcs]
        n, c = len(srcs), srcs[0].shape[0]
        sw, sh = scale_factors([s.shape for s in srcs], ht, wt, scale_ratios)
        hs = np.array([s.shape[1] for s in srcs], np.int32)
        ws = np.array([s.shape[2] for s in srcs], np.int32)
        ptrs = (_f32p * n)(*[_fp(s) for s in srcs])
        out = np.empty((c, ht, wt), np.float32)
        self.lib.refhu_resize_average.restype = None
        self.lib.refhu_resize_average.argtypes = [_f32p, C.POINTER(_f32p), _i32p, _i32p, _f32p, _f32p] + [C.c_int] * 4
        self.lib.refhu_resize_average(_fp(out), ptrs, _ip(hs), _ip(ws), _fp(sh), _fp(sw), n, c, ht, wt)
        return out

    def hu_resize_and_pad(self, planes: np.ndarray, net_size, scale: float) -> np.ndarray:
        """resizeAndPadKernel's pixel walk: planes (3, Hs, Ws) fp32 -> (3, Ht, Wt)."""
        planes = np.ascontiguousarray(planes, np.float32)
        _, hs, ws = planes.shape
        out = np.empty((3, int(net_size[0]), int(net_size[1])), np.float32)
        self.lib.refhu_resize_and_pad.restype = None
        self.lib.refhu_resize_and_pad.argtypes = [_f32p, _f32p] + [C.c_int] * 4 + [C.c_float]
        self.lib.refhu_resize_and_pad(_fp(out), _fp(planes), hs, ws, out.shape[1], out.shape[2], float(scale))
        return out

    def nms(self, heat, thr, channels, off=(0.5, 0.5), max_peaks=MAX_PEAKS, peaks=None) -> np.ndarray:
        heat = np.ascontiguousarray(heat, np.float32)
        _, h, w = heat.shape
        if peaks is None:
            peaks = np.zeros((channels, max_peaks + 1, 3), np.float32)
        kernel = np.zeros((channels, h, w), np.int32)
        tsz = np.array([1, channels, max_peaks + 1, 3], np.int32)
        ssz = np.array([1, heat.shape[0], h, w], np.int32)
        rc = self.lib.ref_nms_cpu(_fp(peaks), _ip(kernel), _fp(heat), thr, _ip(tsz), _ip(ssz), off[0], off[1])
        if rc:
            raise ValueError(self.lib.ref_last_error().decode())
        return peaks

    def connect(self, heat, peaks, model, p=None, scale=1.0, maximize_positives=False, max_peaks=MAX_PEAKS,
                capacity=1024):
        p = p or pose_defaults(model, maximize_positives)
        heat = np.ascontiguousarray(heat, np.float32)
        peaks = np.ascontiguousarray(peaks, np.float32)
        _, h, w = heat.shape
        parts = model_dims(model)[0]
        kp = np.zeros((capacity, parts, 3), np.float32)
        sc = np.zeros((capacity,), np.float32)
        n = self.lib.ref_connect_cpu(_fp(heat), _fp(peaks), model, w, h, max_peaks, p["inter_min_above"],
                                     p["inter_threshold"], int(p["min_subset_cnt"]), p["min_subset_score"],
                                     p["nms_threshold"], scale, int(maximize_positives), _fp(kp), _fp(sc),
                                     capacity)
        if n < 0:
            raise RuntimeError(self.lib.ref_last_error().decode())
        assert n <= capacity
        return kp[:n].copy(), sc[:n].copy()

    def connect_from_pair_scores(self, cube, peaks, model, p=None, scale=1.0, maximize_positives=False,
                                 max_peaks=MAX_PEAKS, capacity=1024):
        p = p or pose_defaults(model, maximize_positives)
        cube = np.ascontiguousarray(cube, np.float32)
        peaks = np.ascontiguousarray(peaks, np.float32)
        parts = model_dims(model)[0]
        kp = np.zeros((capacity, parts, 3), np.float32)
        sc = np.zeros((capacity,), np.float32)
        n = self.lib.ref_connect_from_pair_scores(_fp(cube), _fp(peaks), model, max_peaks,
                                                  int(p["min_subset_cnt"]), p["min_subset_score"], scale,
                                                  int(maximize_positives), _fp(kp), _fp(sc), capacity)
        if n < 0:
            raise RuntimeError(self.lib.ref_last_error().decode())
        return kp[:n].copy(), sc[:n].copy()

    def tables(self, model: int):
        buf = (C.c_uint * 512)()
        n = self.lib.ref_pose_part_pairs(model, buf, 512)
        pairs = list(buf[:n])
        n = self.lib.ref_pose_map_index(model, buf, 512)
        mapidx = list(buf[:n])
        return pairs, mapidx

    def defaults(self, model: int, maximize_positives: bool = False) -> dict:
        out = np.zeros(5, np.float32)
        self.lib.ref_pose_defaults(model, int(maximize_positives), _fp(out))
        return dict(nms_threshold=float(out[0]), inter_min_above=float(out[1]), inter_threshold=float(out[2]),
                    min_subset_cnt=int(out[3]), min_subset_score=float(out[4]))
