"""TEST INFRASTRUCTURE (oracle) -- ctypes front-end of oracle/_ref/libopref_cuda.so: the reference's own CUDA
sources (resizeAndMergeBase.cu, nmsBase.cu, maximumBase.cu, bodyPartConnectorBase.cu, gpu/cuda.cu) compiled
unmodified by oracle/Makefile (`make refcuda`) and driven with torch tensors on the GPU box.

Two uses, both as the checker / the baseline, never as the product:
  * `-m gpu` parity tests: CUDA-build semantics (nms border rule, process() scores, pafVectorIntoPeopleVector,
    FMA-contracted bicubic) straight from the reference's kernels;
  * bench.py: `ref_gpu_ms` per post-processing stage -- the reference's kernels on the same inputs, same box.
The reference launches on the legacy default stream, which is torch's default stream, so torch.cuda.Event timing on
the current (default) stream brackets its launches.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import numpy as np
import torch

from .post import MAX_PEAKS, model_dims, pose_defaults

_HERE = os.path.dirname(os.path.abspath(__file__))
REFCUDA_SO = os.path.join(_HERE, "_ref", "libopref_cuda.so")


def _p(t: torch.Tensor):
    assert t.is_cuda and t.is_contiguous()
    return C.c_void_p(t.data_ptr())


def _i4(shape):
    return (C.c_int * 4)(*[int(v) for v in shape])


class RefCuda:
    @staticmethod
    def available() -> bool:
        return os.path.exists(REFCUDA_SO) and torch.cuda.is_available()

    def __init__(self):
        L = self.lib = C.CDLL(REFCUDA_SO)
        L.refcuda_last_error.restype = C.c_char_p
        v, i, f = C.c_void_p, C.c_int, C.c_float
        L.refcuda_resize_merge.argtypes = [v, C.POINTER(v), C.POINTER(i), C.POINTER(f), i, C.POINTER(i)]
        L.refcuda_nms.argtypes = [v, v, v, f, C.POINTER(i), C.POINTER(i), f, f]
        L.refcuda_maximum.argtypes = [v, v, C.POINTER(i), C.POINTER(i)]
        L.refcuda_connect.argtypes = [v, v, v, i, i, i, i, f, f, i, f, f, f, i, v, v, v, v, v, v, i, C.POINTER(i)]
        L.refcuda_reorder_and_normalize.argtypes = [v, v, i, i, i]
        L.refcuda_resize_and_pad_rbg.argtypes = [v, v, i, i, i, i, f]
        self._tables = {}

    def _check(self, rc):
        if rc:
            raise RuntimeError(self.lib.refcuda_last_error().decode())

    # resizeAndMergeGpu: srcs[i] (N_i, C, h, w) device tensors -> dst (N, C, H, W) (the reference uses frame 0 only
    # for several scales and loops over N for one scale)
    def resize_merge(self, dst: torch.Tensor, srcs: Sequence[torch.Tensor], scale_ratios=None) -> torch.Tensor:
        n = len(srcs)
        ptrs = (C.c_void_p * n)(*[s.data_ptr() for s in srcs])
        shapes = (C.c_int * (4 * n))(*[int(v) for s in srcs for v in s.shape])
        ratios = (C.c_float * n)(*([1.0] * n if scale_ratios is None else [float(r) for r in scale_ratios]))
        self._check(self.lib.refcuda_resize_merge(_p(dst), ptrs, shapes, ratios, n, _i4(dst.shape)))
        return dst

    def nms(self, peaks: torch.Tensor, heat: torch.Tensor, thr: float, off=(0.5, 0.5),
            kernel: Optional[torch.Tensor] = None) -> torch.Tensor:
        if kernel is None:
            # nmsRegisterKernel has no x < w / y < h guard on its border branch (nmsBase.cu:86-87): threads of the
            # 16x16 grid overhang with y == 0 / h-1 or x == 0 / w-1 write zeros at y*w + x, i.e. into the following
            # rows / planes and, for the last plane, past the map -> slack. (For w % 16 != 0 the overhang of row 0
            # lands on row 1's first pixels and RACES with their own threads: callers compare only w % 16 == 0.)
            kernel = torch.empty(heat.numel() + 16 * (heat.shape[-1] + 16), dtype=torch.int32, device=heat.device)
        self._check(self.lib.refcuda_nms(_p(peaks), _p(kernel), _p(heat), thr, _i4(peaks.shape), _i4(heat.shape),
                                         off[0], off[1]))
        return peaks

    def maximum(self, target: torch.Tensor, heat: torch.Tensor) -> torch.Tensor:
        self._check(self.lib.refcuda_maximum(_p(target), _p(heat), _i4(target.shape), _i4(heat.shape)))
        return target

    def _model_tables(self, model: int, device):
        key = (model, str(device))
        if key not in self._tables:
            from .post import Ref
            pairs, mapidx = Ref().tables(model)
            parts = model_dims(model)[0]
            shifted = [m + parts + 1 for m in mapidx]  # bodyPartConnectorCaffe.cpp:353-359 (both models add bkg)
            # torch has no uint32 arithmetic needs here: the bit patterns of small non-negative ints are the same
            self._tables[key] = (torch.tensor(pairs, dtype=torch.int32, device=device),
                                 torch.tensor(shifted, dtype=torch.int32, device=device))
        return self._tables[key]

    # connectBodyPartsGpu for ONE frame: heat (C, H, W) device, peaks (parts, maxPeaks+1, 3) device
    # -> keypoints (P, parts, 3), scores (P), pair-score cube (pairs, maxPeaks, maxPeaks) as numpy
    def connect(self, heat: torch.Tensor, peaks: torch.Tensor, model: int, p=None, scale=1.0,
                maximize_positives=False, max_peaks=MAX_PEAKS, capacity=1024, cube_dev: Optional[torch.Tensor] = None,
                want_cube=True):
        p = p or pose_defaults(model, maximize_positives)
        parts, pairs = model_dims(model)
        _, h, w = heat.shape
        pairs_dev, mapidx_dev = self._model_tables(model, heat.device)
        if cube_dev is None:
            cube_dev = torch.empty((pairs, max_peaks, max_peaks), device=heat.device)
        peaks_host = peaks.cpu().contiguous()  # bottom.at(1)->cpu_data() (bodyPartConnectorCaffe.cpp:341)
        kp = np.zeros((capacity, parts, 3), np.float32)
        sc = np.zeros((capacity,), np.float32)
        cube = np.empty((pairs, max_peaks, max_peaks), np.float32) if want_cube else None
        n = C.c_int(0)
        self._check(self.lib.refcuda_connect(
            _p(heat), C.c_void_p(peaks_host.data_ptr()), _p(peaks), model, w, h, max_peaks, p["inter_min_above"],
            p["inter_threshold"], int(p["min_subset_cnt"]), p["min_subset_score"], p["nms_threshold"], scale,
            int(maximize_positives), _p(cube_dev), _p(pairs_dev), _p(mapidx_dev),
            C.c_void_p(cube.ctypes.data) if want_cube else None, C.c_void_p(kp.ctypes.data),
            C.c_void_p(sc.ctypes.data), capacity, C.byref(n)))
        m = min(n.value, capacity)
        return kp[:m].copy(), sc[:m].copy(), cube

    # reorderAndNormalize + resizeAndPadRbgGpu = the device branch of CvMatToOpInput::createArray
    # (core/cvMatToOpInput.cpp:96-137): frame (Hs, Ws, 3) u8 device -> (3, Ht, Wt) fp32 device
    def preprocess(self, frame: torch.Tensor, net_size, scale: float) -> torch.Tensor:
        hs, ws, ch = frame.shape
        planes = torch.empty((ch, hs, ws), device=frame.device)
        self._check(self.lib.refcuda_reorder_and_normalize(_p(planes), _p(frame), ws, hs, ch))
        out = torch.empty((3, int(net_size[0]), int(net_size[1])), device=frame.device)
        self._check(self.lib.refcuda_resize_and_pad_rbg(_p(out), _p(planes), ws, hs, out.shape[2], out.shape[1],
                                                        float(scale)))
        return out
