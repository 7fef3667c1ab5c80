"""Host-side mirror of the reference's hand / face keypoint extractors for the hot path:
crop -> net -> resizeAndMerge x8 -> maximum -> affine back-projection.

Mirrors (reference files relative to zengjianping/openpose, same argument meaning / error text):
  HandExtractorB200.forwardPass <- HandExtractorCaffe::forwardPass  src/openpose/hand/handExtractorCaffe.cpp:305-464
      crop_matrix + crop_frames_gpu <- cropFrame                    handExtractorCaffe.cpp:44-76
      connect_keypoints         <- connectKeypoints                 handExtractorCaffe.cpp:78-102
  FaceExtractorB200.forwardPass <- FaceExtractorCaffe::forwardPass  src/openpose/face/faceExtractorCaffe.cpp:175-300
  get_average_score             <- getAverageScore                  src/openpose/utilities/keypoint.cpp:352-373
  u_char_to_float               <- uCharCvMatToFloatPtr(normalize=1) src/openpose/utilities/openCv.cpp:57-130

B200-first difference (results per crop are the reference's): the reference runs its net once per crop, batch 1,
one after the other; here every valid crop of the frame (both hands of every person, every scale) goes through
ONE batched net forward, one resize launch and one maximum launch -- the unit SURVEY 8(e) shards across GPUs.
The frame is uploaded once as u8 and the crops themselves (cv::warpAffine INTER_LINEAR | WARP_INVERSE_MAP +
uCharCvMatToFloatPtr in the reference, on the host, one crop at a time) are cut on the device by one launch of
opb200_crop_warp, bit-identical to OpenCV's 8-bit fixed-point result.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib
from . import net as netmod
from . import post
from ._lib import Opb200Error
from .pose_extractor import HeatMapType, ScaleMode, maximum_top_shape, positive_int_round, \
    resize_and_merge_top_shape

HAND_NUMBER_PARTS = 21           # include/openpose/hand/handParameters.hpp:11
HAND_CCN_DECREASE_FACTOR = 8.0   # handParameters.hpp:43
FACE_NUMBER_PARTS = 70           # include/openpose/face/faceParameters.hpp:11
FACE_CCN_DECREASE_FACTOR = 8.0   # faceParameters.hpp:24

Rect = Tuple[float, float, float, float]   # x, y, width, height (op::Rectangle<float>)


def u_char_to_float(image_hwc: np.ndarray) -> np.ndarray:
    """uint8 H x W x C -> float32 C x H x W, x/256 - 0.5 (exact in float: the AVX fmadd and the scalar path agree)."""
    return (image_hwc.transpose(2, 0, 1).astype(np.float32) * np.float32(1.0 / 256.0)) - np.float32(0.5)


def crop_matrix(rect: Rect, net_input_side: int, mirror: bool) -> np.ndarray:
    """The 2x3 CV_64F inverse map cropFrame builds (handExtractorCaffe.cpp:53-68): crop pixel (u,v) <- image
    (a00*u + a02, a11*v + a12); float arithmetic for the scale and the mirrored offset, as the reference."""
    scale = np.float32(rect[2]) / np.float32(net_input_side)
    m = np.eye(2, 3, dtype=np.float64)
    m[0, 0] = -float(scale) if mirror else float(scale)
    m[1, 1] = float(scale)
    m[0, 2] = float(np.float32(rect[0]) + np.float32(rect[2])) if mirror else float(np.float32(rect[0]))
    m[1, 2] = float(np.float32(rect[1]))
    return m


def upload_frame(cv_input, device) -> torch.Tensor:
    """(H,W,3) uint8 BGR frame (numpy or torch, host or device) -> contiguous CUDA tensor; one H2D per frame."""
    if cv_input is None or (cv_input.numel() if isinstance(cv_input, torch.Tensor) else cv_input.size) == 0:
        raise Opb200Error("Empty cvInputData.")
    t = cv_input if isinstance(cv_input, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(cv_input))
    if t.dim() != 3 or t.shape[2] != 3 or t.dtype != torch.uint8:
        raise Opb200Error("Input images must be 3-channel BGR.")
    return t.to(device, non_blocking=True).contiguous()


def crop_frames_gpu(frame: torch.Tensor, matrices: Sequence[np.ndarray], net_output_size: Tuple[int, int]) -> torch.Tensor:
    """All crops of one frame in one launch: cv::warpAffine(INTER_LINEAR | WARP_INVERSE_MAP, BORDER_CONSTANT 0) +
    uCharCvMatToFloatPtr (cropFrame, handExtractorCaffe.cpp:44-76; faceExtractorCaffe.cpp:133-149) ->
    (N,3,H,W) float32 on the frame's device (libopb200: opb200_crop_warp)."""
    mats = torch.from_numpy(np.ascontiguousarray(np.stack(matrices), np.float64).reshape(-1, 6)).to(frame.device)
    out = torch.empty((mats.shape[0], 3, int(net_output_size[1]), int(net_output_size[0])), dtype=torch.float32,
                      device=frame.device)
    L = _lib.lib()
    _lib.check(L.opb200_crop_warp(out.data_ptr(), frame.data_ptr(), mats.data_ptr(), mats.shape[0], frame.shape[0],
                                  frame.shape[1], out.shape[2], out.shape[3],
                                  torch.cuda.current_stream(frame.device).cuda_stream))
    return out


def connect_keypoints(peaks: np.ndarray, affine: np.ndarray) -> np.ndarray:
    """peaks (parts, 3) in crop pixels -> (parts, 3) in image pixels, double-precision affine then float."""
    x, y = peaks[:, 0].astype(np.float64), peaks[:, 1].astype(np.float64)
    out = np.empty_like(peaks, dtype=np.float32)
    out[:, 0] = (affine[0, 0] * x + affine[0, 1] * y + affine[0, 2]).astype(np.float32)
    out[:, 1] = (affine[1, 0] * x + affine[1, 1] * y + affine[1, 2]).astype(np.float32)
    out[:, 2] = peaks[:, 2]
    return out


def get_average_score(keypoints: np.ndarray) -> np.float32:
    """float accumulation in part order, then / numberKeypoints."""
    s = np.float32(0)
    for v in keypoints[:, 2]:
        s = np.float32(s + np.float32(v))
    return np.float32(s / np.float32(keypoints.shape[0]))


def recenter(rect: Rect, new_w: float, new_h: float) -> Rect:
    f = np.float32
    cx = f(f(rect[0]) + f(rect[2]) / f(2))
    cy = f(f(rect[1]) + f(rect[3]) / f(2))
    return (float(cx - f(f(new_w) / f(2.0))), float(cy - f(f(new_h) / f(2.0))), float(new_w), float(new_h))


class _CropNet:
    """Net + resize + maximum on a batch of crops (what detectHandKeypoints does per crop)."""

    def __init__(self, proto: str, model: str, gpu: int, decrease_factor: float):
        self.net = netmod.NetB200(proto, model, gpuId=gpu)
        self.gpu, self.factor = gpu, decrease_factor
        self.heat: Optional[torch.Tensor] = None
        self.peaks: Optional[torch.Tensor] = None

    def init(self) -> None:
        self.net.initializationOnThread()

    def run(self, crops) -> np.ndarray:
        """crops (N,3,H,W) float32 (CUDA tensor, or numpy) -> peaks (N, parts, 3) in crop pixels (host)."""
        self.net.forwardPass(crops if isinstance(crops, torch.Tensor) else np.ascontiguousarray(crops, np.float32))
        out = self.net.getOutputBlobArray()
        top = resize_and_merge_top_shape(list(out.shape), self.factor, 1.0, False)
        if self.heat is None or list(self.heat.shape) != top:
            dev = torch.device("cuda", self.gpu)
            self.heat = torch.empty(top, dtype=torch.float32, device=dev)
            self.peaks = torch.empty(maximum_top_shape(top), dtype=torch.float32, device=dev)
        post.resize_and_merge_gpu(self.heat, [out])
        post.maximum_gpu(self.peaks, self.heat)
        return self.peaks[:, 0].cpu().numpy()


def _heat_maps_for_crops(heat: torch.Tensor, parts: int, mode: ScaleMode) -> np.ndarray:
    """updateHandHeatMapsForPerson / updateFaceHeatMapsForPerson for every crop at once: the part channels under the
    body-part remap rules, written by one kernel straight into pinned host memory -> (N, parts, H, W)."""
    with torch.cuda.device(heat.device):
        return post.heatmaps_export_gpu(heat, parts, True, 0, post.HEATMAP_PARTS, mode).numpy()


class HandExtractorB200:
    def __init__(self, netInputSize: Tuple[int, int], netOutputSize: Tuple[int, int], caffeProto: str,
                 caffeTrainedModel: str, gpuId: int = 0, numberScales: int = 1, rangeScales: float = 0.4,
                 heatMapTypes: Sequence[HeatMapType] = (), heatMapScaleMode: ScaleMode = ScaleMode.ZeroToOneFixedAspect):
        if tuple(netInputSize) != tuple(netOutputSize):      # handExtractorNet.cpp ctor check
            raise Opb200Error("Net input and output size must be equal.")
        self.mNetOutputSize = (int(netOutputSize[0]), int(netOutputSize[1]))
        self.mMultiScaleNumberAndRange = (int(numberScales), float(rangeScales))
        self.mHeatMapTypes, self.mHeatMapScaleMode = list(heatMapTypes), ScaleMode(heatMapScaleMode)
        self._net = _CropNet(caffeProto, caffeTrainedModel, gpuId, HAND_CCN_DECREASE_FACTOR)
        self.mHandKeypoints = [np.zeros((0, HAND_NUMBER_PARTS, 3), np.float32) for _ in range(2)]
        self.mHeatMaps: List[Optional[np.ndarray]] = [None, None]
        self.mEnabled = True

    def initializationOnThread(self) -> None:
        self._net.init()

    def getHandKeypoints(self) -> List[np.ndarray]:
        return self.mHandKeypoints

    def getHeatMaps(self) -> List[Optional[np.ndarray]]:
        return self.mHeatMaps

    def forwardPass(self, handRectangles: Sequence[Sequence[Rect]], cvInputData: np.ndarray) -> None:
        """handRectangles[person] = (left, right) rectangles; cvInputData: H x W x 3 uint8 BGR frame."""
        if not self.mEnabled or len(handRectangles) == 0:
            self.mHandKeypoints = [np.zeros((0, HAND_NUMBER_PARTS, 3), np.float32) for _ in range(2)]
            return
        if cvInputData is None or cvInputData.size == 0:
            raise Opb200Error("Empty cvInputData.")
        side = min(self.mNetOutputSize)
        people = len(handRectangles)
        self.mHandKeypoints = [np.zeros((people, HAND_NUMBER_PARTS, 3), np.float32) for _ in range(2)]
        want_heat = bool(self.mHeatMapTypes)
        if want_heat:
            self.mHeatMaps = [np.zeros((people, HAND_NUMBER_PARTS, self.mNetOutputSize[1], self.mNetOutputSize[0]),
                                       np.float32) for _ in range(2)]
        n_scales, rng = self.mMultiScaleNumberAndRange
        jobs = []                     # job = (hand, person, scale index, affine)
        for hand in range(2):
            mirror = hand == 0
            for person in range(people):
                r = tuple(float(v) for v in handRectangles[person][hand])
                if r[2] != r[3]:
                    raise Opb200Error("Hand rectangle for hand keypoint estimation must be squared, i.e., width = "
                                      "height (%f vs. %f)." % (r[2], r[3]))
                if not (min(r[2], r[3]) > 1 and r[2] * r[3] > 10):
                    continue
                if n_scales == 1:
                    scaled = [r]
                else:
                    init = np.float32(1.0) - np.float32(rng) / np.float32(2.0)
                    scaled = []
                    for i in range(n_scales):
                        s = init + np.float32(rng) * np.float32(i) / (np.float32(n_scales) - np.float32(1.0))
                        w = float(positive_int_round(float(np.float32(r[2]) * s)) // 2 * 2)
                        h = float(positive_int_round(float(np.float32(r[3]) * s)) // 2 * 2)
                        scaled.append(recenter(r, w, h))
                for i, rs in enumerate(scaled):
                    jobs.append((hand, person, i, crop_matrix(rs, side, mirror)))
        if not jobs:
            return
        frame = upload_frame(cvInputData, torch.device("cuda", self._net.gpu))
        peaks = self._net.run(crop_frames_gpu(frame, [j[3] for j in jobs], self.mNetOutputSize))
        best = {}
        for j, (hand, person, i, aff) in enumerate(jobs):
            est = connect_keypoints(peaks[j], aff)
            cur = self.mHandKeypoints[hand][person]
            if i == 0 or get_average_score(est) > get_average_score(cur):
                self.mHandKeypoints[hand][person] = est
            best[(hand, person)] = j        # the reference stores the heat-maps of the LAST scale processed
        if want_heat:
            exported = _heat_maps_for_crops(self._net.heat, HAND_NUMBER_PARTS, self.mHeatMapScaleMode)
            for (hand, person), j in best.items():
                self.mHeatMaps[hand][person] = exported[j]


class FaceExtractorB200:
    def __init__(self, netInputSize: Tuple[int, int], netOutputSize: Tuple[int, int], caffeProto: str,
                 caffeTrainedModel: str, gpuId: int = 0, heatMapTypes: Sequence[HeatMapType] = (),
                 heatMapScaleMode: ScaleMode = ScaleMode.ZeroToOneFixedAspect):
        if tuple(netInputSize) != tuple(netOutputSize):
            raise Opb200Error("Net input and output size must be equal.")
        self.mNetOutputSize = (int(netOutputSize[0]), int(netOutputSize[1]))
        self.mHeatMapTypes, self.mHeatMapScaleMode = list(heatMapTypes), ScaleMode(heatMapScaleMode)
        self._net = _CropNet(caffeProto, caffeTrainedModel, gpuId, FACE_CCN_DECREASE_FACTOR)
        self.mFaceKeypoints = np.zeros((0, FACE_NUMBER_PARTS, 3), np.float32)
        self.mHeatMaps: Optional[np.ndarray] = None
        self.mEnabled = True

    def initializationOnThread(self) -> None:
        self._net.init()

    def getFaceKeypoints(self) -> np.ndarray:
        return self.mFaceKeypoints

    def getHeatMaps(self) -> Optional[np.ndarray]:
        return self.mHeatMaps

    def forwardPass(self, faceRectangles: Sequence[Rect], cvInputData: np.ndarray) -> None:
        if not self.mEnabled or len(faceRectangles) == 0:
            self.mFaceKeypoints = np.zeros((0, FACE_NUMBER_PARTS, 3), np.float32)
            return
        if cvInputData is None or cvInputData.size == 0:
            raise Opb200Error("Empty cvInputData.")
        side = min(self.mNetOutputSize)
        people = len(faceRectangles)
        self.mFaceKeypoints = np.zeros((people, FACE_NUMBER_PARTS, 3), np.float32)
        want_heat = bool(self.mHeatMapTypes)
        if want_heat:
            self.mHeatMaps = np.zeros((people, FACE_NUMBER_PARTS, self.mNetOutputSize[1], self.mNetOutputSize[0]),
                                      np.float32)
        jobs = []
        for person in range(people):
            r = tuple(float(v) for v in faceRectangles[person])
            if r[2] != r[3]:
                raise Opb200Error("Face rectangle for face keypoint estimation must be squared, i.e., width = "
                                  "height (%f vs. %f)." % (r[2], r[3]))
            if not min(r[2], r[3]) > 40:          # only faces with a minimum pixel area
                continue
            scale_face = float(np.float32(max(r[2], r[3]))) / float(side)     # float / double
            m = np.eye(2, 3, dtype=np.float64)
            m[0, 0] = m[1, 1] = scale_face
            m[0, 2], m[1, 2] = float(np.float32(r[0])), float(np.float32(r[1]))
            jobs.append((person, m))
        if not jobs:
            return
        frame = upload_frame(cvInputData, torch.device("cuda", self._net.gpu))
        peaks = self._net.run(crop_frames_gpu(frame, [j[1] for j in jobs], self.mNetOutputSize))
        exported = _heat_maps_for_crops(self._net.heat, FACE_NUMBER_PARTS, self.mHeatMapScaleMode) if want_heat else None
        for j, (person, m) in enumerate(jobs):
            self.mFaceKeypoints[person] = connect_keypoints(peaks[j], m)
            if want_heat:
                self.mHeatMaps[person] = exported[j]
