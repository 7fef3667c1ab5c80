"""Frame -> net input: the callers upstream of op::Net::forwardPass, device-resident.

ScaleAndSizeExtractor mirrors src/openpose/core/scaleAndSizeExtractor.cpp:7-117 (host arithmetic only).
CvMatToOpInputB200 mirrors the device branch of op::CvMatToOpInput::createArray
(src/openpose/core/cvMatToOpInput.cpp:61-161): same arguments, same error texts, but frames stay on the GPU and a
batch of frames goes through one launch of libopb200's opb200_preprocess (reorderAndNormalize + resizeAndPadRbgGpu
fused, csrc/post/preprocess.cu). There is no CPU path here: without the CUDA library every call raises.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np
import torch

from . import _lib
from ._lib import Opb200Error
from .pose_extractor import positive_int_round, resize_get_scale_factor

Point = Tuple[int, int]  # (x, y), as op::Point<int>


def _f32(v: float) -> float:
    return float(np.float32(v))


class ScaleAndSizeExtractor:
    """reference: include/openpose/core/scaleAndSizeExtractor.hpp, scaleAndSizeExtractor.cpp:7-31."""

    def __init__(self, netInputResolution: Point, netInputResolutionDynamicBehavior: float = -1.0,
                 outputResolution: Point = (-1, -1), scaleNumber: int = 1, scaleGap: float = 0.25):
        nx, ny = netInputResolution
        if (nx > 0 and nx % 16 != 0) or (ny > 0 and ny % 16 != 0):
            raise Opb200Error("Net input resolution must be multiples of 16.")
        if scaleNumber < 1:
            raise Opb200Error("There must be at least 1 scale.")
        if scaleGap <= 0.0:
            raise Opb200Error("The gap between scales must be strictly positive.")
        self.mNetInputResolution = (int(nx), int(ny))
        self.mNetInputResolutionDynamicBehavior = float(netInputResolutionDynamicBehavior)
        self.mOutputSize = (int(outputResolution[0]), int(outputResolution[1]))
        self.mScaleNumber = int(scaleNumber)
        self.mScaleGap = float(scaleGap)

    def extract(self, inputResolution: Point):
        """-> (scaleInputToNetInputs, netInputSizes, scaleInputToOutput, outputResolution); .cpp:37-117."""
        ix, iy = int(inputResolution[0]), int(inputResolution[1])
        if ix * iy <= 0:
            raise Opb200Error("Wrong input element (empty cvInputData).")
        px, py = self.mNetInputResolution
        if px <= 0 or py <= 0:
            if px <= 0 and py <= 0:
                raise Opb200Error("Only 1 of the dimensions of net input resolution can be <= 0.")
            dyn = self.mNetInputResolutionDynamicBehavior
            # float arithmetic as in the reference (:54-69)
            if dyn > 0:
                if px <= 0:
                    a = _f32(_f32(_f32(py * _f32(dyn)) * 16.0) / 9.0)
                    b = _f32(_f32(py * ix) / _f32(iy))
                    px = 16 * positive_int_round(_f32(_f32(1 / 16.0) * min(a, b)))
                else:
                    a = _f32(_f32(_f32(px * _f32(dyn)) * 9.0) / 16.0)
                    b = _f32(_f32(px * iy) / _f32(ix))
                    py = 16 * positive_int_round(_f32(_f32(1 / 16.0) * min(a, b)))
            else:
                if px <= 0:
                    px = 16 * positive_int_round(_f32(_f32(_f32(_f32(1 / 16.0) * py) * ix) / _f32(iy)))
                else:
                    py = 16 * positive_int_round(_f32(_f32(_f32(_f32(1 / 16.0) * px) * iy) / _f32(ix)))
        scales: List[float] = []
        sizes: List[Point] = []
        for i in range(self.mScaleNumber):
            cur = 1.0 - i * self.mScaleGap
            if cur < 0.0 or 1.0 < cur:
                raise Opb200Error("All scales must be in the range [0, 1], i.e., 0 <= 1-scale_number*scale_gap <= 1")
            tw = min(px, max(1, positive_int_round(px * cur) // 16 * 16))
            th = min(py, max(1, positive_int_round(py * cur) // 16 * 16))
            scales.append(resize_get_scale_factor((ix, iy), (tw, th)))
            sizes.append((tw, th))
        if self.mOutputSize[0] > 0 and self.mOutputSize[1] > 0:
            out_res = self.mOutputSize
            scale_out = resize_get_scale_factor((ix, iy), out_res)
        else:
            out_res = (ix, iy)
            scale_out = 1.0
        return scales, sizes, scale_out, out_res


def preprocess_gpu(dst: torch.Tensor, src: torch.Tensor, scale: float) -> torch.Tensor:
    """dst (N,3,Ht,Wt) fp32 <- src (N,Hs,Ws,3) u8, both on the device; one launch."""
    if not (src.is_cuda and src.dtype == torch.uint8 and src.is_contiguous() and src.dim() == 4):
        raise Opb200Error("src must be a contiguous (N,Hs,Ws,3) uint8 CUDA tensor")
    if src.shape[3] != 3:
        raise Opb200Error("Input images must be 3-channel BGR.")
    if not (dst.is_cuda and dst.dtype == torch.float32 and dst.is_contiguous() and dst.dim() == 4
            and dst.shape[0] == src.shape[0] and dst.shape[1] == 3):
        raise Opb200Error("dst must be a contiguous (N,3,Ht,Wt) float32 CUDA tensor")
    L = _lib.lib()
    _lib.check(L.opb200_preprocess(dst.data_ptr(), src.data_ptr(), src.shape[0], src.shape[1], src.shape[2],
                                   dst.shape[2], dst.shape[3], float(scale),
                                   torch.cuda.current_stream().cuda_stream))
    return dst


class CvMatToOpInputB200:
    """reference: include/openpose/core/cvMatToOpInput.hpp; createArray = cvMatToOpInput.cpp:61-161."""

    def __init__(self, poseModel: int = 0, gpuResize: bool = True, device: str = "cuda:0"):
        if not gpuResize:
            # The reference's CPU branch is OpenCV (cv::warpAffine); this package has no host compute path.
            raise Opb200Error("CvMatToOpInputB200 only has the device path (gpuResize = true).")
        self.mPoseModel = int(poseModel)
        self.device = torch.device(device)
        _lib.lib()

    def createArray(self, inputData, scaleInputToNetInputs: Sequence[float],
                    netInputSizes: Sequence[Point]) -> List[torch.Tensor]:
        """inputData: (H,W,3) or (N,H,W,3) uint8 BGR, numpy or torch (host or device).
        Returns one (N,3,netH,netW) fp32 device tensor per scale (N = 1 for a single frame, as the reference)."""
        frames = inputData if isinstance(inputData, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(inputData))
        if frames.numel() == 0:
            raise Opb200Error("Wrong input element (empty inputData).")
        if frames.dim() == 3:
            frames = frames.unsqueeze(0)
        if frames.dim() != 4 or frames.shape[3] != 3:
            raise Opb200Error("Input images must be 3-channel BGR.")
        if len(scaleInputToNetInputs) != len(netInputSizes):
            raise Opb200Error("scaleInputToNetInputs.size() != netInputSizes.size().")
        if frames.dtype != torch.uint8:
            raise Opb200Error("Input images must be 8-bit.")
        frames = frames.to(self.device, non_blocking=True).contiguous()
        out = []
        for scale, (nx, ny) in zip(scaleInputToNetInputs, netInputSizes):
            dst = torch.empty((frames.shape[0], 3, int(ny), int(nx)), dtype=torch.float32, device=self.device)
            out.append(preprocess_gpu(dst, frames, scale))
        return out
