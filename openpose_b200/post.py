"""Python mirror of the reference's post-processing operator interface, calling the CUDA kernels through
the C ABI (include/opb200.h). torch is used only for device memory and streams.

Mirrors (reference files relative to zengjianping/openpose):
  resize_and_merge_gpu  <- op::resizeAndMergeGpu   include/openpose/net/resizeAndMergeBase.hpp:14-17
  nms_gpu               <- op::nmsGpu              include/openpose/net/nmsBase.hpp:14-17
  maximum_gpu           <- op::maximumGpu          include/openpose/net/maximumBase.hpp
  connect_body_parts_gpu<- op::connectBodyPartsGpu include/openpose/net/bodyPartConnectorBase.hpp:17-24
Argument meaning and error behaviour follow those functions (errors raise Opb200Error with the reference's
message text where the reference has one).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import ConnectParams, Opb200Error, POSE_BODY_25, POSE_COCO_18, SEMANTICS_CPU, SEMANTICS_CUDA

POSE_MAX_PEOPLE = 127  # reference: include/openpose/pose/poseParameters.hpp:14


def _stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def _dev(t: torch.Tensor, name: str) -> int:
    if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
        raise Opb200Error("%s must be a contiguous float32 CUDA tensor" % name)
    return t.data_ptr()


_scratch: Dict[Tuple[str, int], torch.Tensor] = {}


def _get_scratch(kind: str, nbytes: int, device) -> torch.Tensor:
    key = (kind, torch.device(device).index or 0)
    buf = _scratch.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.zeros(max(nbytes, 1024), dtype=torch.uint8, device=device)
        _scratch[key] = buf
    return buf


def resize_and_merge_gpu(target: torch.Tensor, sources: Sequence[torch.Tensor],
                         scale_input_to_net_inputs: Optional[Sequence[float]] = None) -> torch.Tensor:
    """target (N,C,H,W) <- average over scales of bicubic-upsampled sources[i] (N,C,h_i,w_i)."""
    L = _lib.lib()
    n = len(sources)
    if n == 0:
        raise Opb200Error("sourceSizes cannot be empty.")
    ptrs = (C.c_void_p * n)(*[_dev(s, "source") for s in sources])
    shapes = (_lib._I4 * n)(*[_lib.i4(s.shape) for s in sources])
    ratios = (C.c_float * n)(*([1.0] * n if scale_input_to_net_inputs is None else
                               [float(r) for r in scale_input_to_net_inputs]))
    _lib.check(L.opb200_resize_merge(_dev(target, "target"), ptrs, shapes, ratios, n, _lib.i4(target.shape),
                                     _stream_ptr()))
    return target


def nms_gpu(peaks: torch.Tensor, heat: torch.Tensor, threshold: float, offset=(0.5, 0.5),
            mode: int = SEMANTICS_CPU) -> torch.Tensor:
    """peaks (N,parts,maxPeaks+1,3) <- heat (N,Ch,H,W); rows beyond each count are left untouched."""
    L = _lib.lib()
    ps, hs = _lib.i4(peaks.shape), _lib.i4(heat.shape)
    scratch = _get_scratch("nms", L.opb200_nms_scratch_bytes(ps, hs), heat.device)
    _lib.check(L.opb200_nms(_dev(peaks, "peaks"), scratch.data_ptr(), _dev(heat, "heat"), float(threshold),
                            float(offset[0]), float(offset[1]), ps, hs, int(mode), _stream_ptr()))
    return peaks


def maximum_gpu(target: torch.Tensor, source: torch.Tensor) -> torch.Tensor:
    """target (N,1,C-1,3) <- per-channel first argmax of source (N,C,H,W)."""
    L = _lib.lib()
    ts, ss = _lib.i4(target.shape), _lib.i4(source.shape)
    scratch = _get_scratch("max", L.opb200_maximum_scratch_bytes(ts, ss), source.device)
    _lib.check(L.opb200_maximum(_dev(target, "target"), scratch.data_ptr(), _dev(source, "source"), ts, ss,
                                _stream_ptr()))
    return target


HEATMAP_PARTS, HEATMAP_BACKGROUND, HEATMAP_PAFS = 1, 2, 4      # 1 << op::HeatMapType


def heatmaps_export_gpu(heat: torch.Tensor, parts: int, has_bkg: bool, paf_channels: int, types_mask: int,
                        scale_mode: int, to_host: bool = True) -> torch.Tensor:
    """getHeatMapsCopy's group selection + range remap (poseExtractorNet.cpp:106-243) for a batch (N,C,H,W) in one
    launch. to_host: the kernel writes straight into pinned host memory (returned tensor, valid after the sync done
    here); otherwise a device tensor."""
    L = _lib.lib()
    n, _, h, w = heat.shape
    ch = L.opb200_heatmaps_export_channels(parts, int(has_bkg), paf_channels, types_mask)
    if to_host:
        dst = torch.empty((n, ch, h, w), dtype=torch.float32, pin_memory=True)
    else:
        dst = torch.empty((n, ch, h, w), dtype=torch.float32, device=heat.device)
    _lib.check(L.opb200_heatmaps_export(dst.data_ptr(), _dev(heat, "heat"), _lib.i4(heat.shape), parts, int(has_bkg),
                                        paf_channels, types_mask, int(scale_mode), _stream_ptr()))
    if to_host:
        torch.cuda.current_stream().synchronize()
    return dst


def pose_defaults(pose_model: int, maximize_positives: bool = False) -> dict:
    """Default thresholds, reference: src/openpose/pose/poseParameters.cpp:677-756 (BODY_25 / COCO_18)."""
    mp = maximize_positives
    return dict(nms_threshold=0.02 if mp else 0.05, inter_min_above=0.75 if mp else 0.95,
                inter_threshold=0.01 if mp else 0.05, min_subset_cnt=2 if mp else 3,
                min_subset_score=0.05 if mp else 0.4)


def make_connect_params(pose_model: int, scale_factor: float = 1.0, maximize_positives: bool = False,
                        semantics: int = SEMANTICS_CPU, max_peaks: int = POSE_MAX_PEOPLE,
                        overrides: Optional[dict] = None) -> ConnectParams:
    d = pose_defaults(pose_model, maximize_positives)
    if overrides:
        d.update(overrides)
    return ConnectParams(pose_model, max_peaks, d["inter_min_above"], d["inter_threshold"],
                         int(d["min_subset_cnt"]), d["min_subset_score"], d["nms_threshold"], scale_factor,
                         int(maximize_positives), semantics)


def model_dims(pose_model: int) -> Tuple[int, int]:
    return {POSE_BODY_25: (25, 26), POSE_COCO_18: (18, 19)}[pose_model]


def connect_body_parts_gpu(heat: torch.Tensor, peaks: torch.Tensor, params: ConnectParams,
                           capacity: int = 512, out=None):
    """Device-side bodyPartConnector. Returns (keypoints (N,capacity,parts,3), scores (N,capacity),
    counts (N,) int32) device tensors; only [:counts[n]] rows of frame n are meaningful."""
    L = _lib.lib()
    n = heat.shape[0]
    parts, _ = model_dims(params.pose_model)
    if out is None:
        kp = torch.empty((n, capacity, parts, 3), dtype=torch.float32, device=heat.device)
        sc = torch.empty((n, capacity), dtype=torch.float32, device=heat.device)
        cnt = torch.empty((n,), dtype=torch.int32, device=heat.device)
    else:
        kp, sc, cnt = out
    scratch = _get_scratch("connect", L.opb200_connect_scratch_bytes(C.byref(params), n), heat.device)
    _lib.check(L.opb200_connect(kp.data_ptr(), sc.data_ptr(), cnt.data_ptr(), capacity, scratch.data_ptr(),
                                _dev(heat, "heat"), _dev(peaks, "peaks"), _lib.i4(heat.shape), C.byref(params),
                                _stream_ptr()))
    return kp, sc, cnt


def paf_scores_gpu(heat: torch.Tensor, peaks: torch.Tensor, params: ConnectParams) -> torch.Tensor:
    """(N, pairs, maxPeaks, maxPeaks) score cube, -1 outside nA x nB (pafScoreKernel layout)."""
    L = _lib.lib()
    n = heat.shape[0]
    _, pairs = model_dims(params.pose_model)
    cube = torch.empty((n, pairs, params.max_peaks, params.max_peaks), dtype=torch.float32, device=heat.device)
    _lib.check(L.opb200_paf_scores(cube.data_ptr(), _dev(heat, "heat"), _dev(peaks, "peaks"),
                                   _lib.i4(heat.shape), C.byref(params), _stream_ptr()))
    return cube
