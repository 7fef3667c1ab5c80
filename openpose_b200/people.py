"""People post-steps on the extractor boundary, device-resident (libopb200: csrc/post/people.cu).

Mirrors (reference files relative to zengjianping/openpose; same names and argument meaning):
  KeepTopNPeople.keepTopPeople  <- src/openpose/core/keepTopNPeople.cpp:17-91
  KeypointScaler.scale          <- src/openpose/core/keypointScaler.cpp:44-104 (+ getScaleAndOffset :7-42)
  HandDetector.detectHands      <- src/openpose/hand/handDetector.cpp:136-164
  FaceDetector.detectFaces      <- src/openpose/face/faceDetector.cpp:120-141
All of them take the connector's batched device layout: keypoints (N,capacity,parts,3), scores (N,capacity),
count (N) int32 CUDA tensors -- what post.connect_body_parts_gpu returns -- and launch one kernel for the batch.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import Opb200Error


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _check_people(keypoints: torch.Tensor, count: Optional[torch.Tensor]) -> Tuple[int, int, int]:
    if not (keypoints.is_cuda and keypoints.dtype == torch.float32 and keypoints.is_contiguous()
            and keypoints.dim() == 4 and keypoints.shape[3] == 3):
        raise Opb200Error("keypoints must be a contiguous (N,capacity,parts,3) float32 CUDA tensor")
    if count is not None and not (count.is_cuda and count.dtype == torch.int32 and count.numel() == keypoints.shape[0]):
        raise Opb200Error("count must be an (N) int32 CUDA tensor")
    return keypoints.shape[0], keypoints.shape[1], keypoints.shape[2]


class KeepTopNPeople:
    def __init__(self, numberPeopleMax: int):
        self.mNumberPeopleMax = int(numberPeopleMax)

    def keepTopPeople(self, keypoints: torch.Tensor, scores: torch.Tensor, count: torch.Tensor):
        """-> (keypoints, scores, count) with at most numberPeopleMax people per frame, original order kept."""
        n, cap, parts = _check_people(keypoints, count)
        out_kp, out_sc, out_n = torch.zeros_like(keypoints), torch.zeros_like(scores), torch.zeros_like(count)
        L = _lib.lib()
        with torch.cuda.device(keypoints.device):
            _lib.check(L.opb200_keep_top_people(out_kp.data_ptr(), out_sc.data_ptr(), out_n.data_ptr(),
                                                keypoints.data_ptr(), scores.data_ptr(), count.data_ptr(), n, cap,
                                                parts, self.mNumberPeopleMax, _stream()))
        return out_kp, out_sc, out_n


class KeypointScaler:
    def __init__(self, scaleMode: int):
        self.mScaleMode = int(scaleMode)

    def scaleAndOffset(self, scaleInputToOutput: float, scaleNetToOutput: float, producerSize) -> Tuple[float, ...]:
        out = (C.c_float * 4)()
        _lib.check(_lib.lib().opb200_keypoint_scale_and_offset(out, self.mScaleMode, float(scaleInputToOutput),
                                                               float(scaleNetToOutput), int(producerSize[0]),
                                                               int(producerSize[1])))
        return tuple(out)

    def scale(self, keypoints: torch.Tensor, scaleInputToOutput: float, scaleNetToOutput: float, producerSize,
              count: Optional[torch.Tensor] = None) -> torch.Tensor:
        """In place, like the reference. keypoints (N,capacity,parts,3); count None = every row."""
        if self.mScaleMode == 0:          # ScaleMode::InputResolution: untouched (keypointScaler.cpp:82)
            return keypoints
        n, cap, parts = _check_people(keypoints, count)
        ox, oy, sx, sy = self.scaleAndOffset(scaleInputToOutput, scaleNetToOutput, producerSize)
        with torch.cuda.device(keypoints.device):
            _lib.check(_lib.lib().opb200_scale_keypoints(keypoints.data_ptr(), count.data_ptr() if count is not None else None,
                                                         n, cap, parts, sx, sy, ox, oy, _stream()))
        return keypoints


def hand_face_rectangles_gpu(keypoints: torch.Tensor, count: torch.Tensor, pose_model: int, hands: bool = True,
                             faces: bool = True):
    """-> (hand rectangles (N,capacity,2,4) or None, face rectangles (N,capacity,4) or None), rows >= count are 0."""
    n, cap, _ = _check_people(keypoints, count)
    h = torch.zeros((n, cap, 2, 4), dtype=torch.float32, device=keypoints.device) if hands else None
    f = torch.zeros((n, cap, 4), dtype=torch.float32, device=keypoints.device) if faces else None
    with torch.cuda.device(keypoints.device):
        _lib.check(_lib.lib().opb200_hand_face_rectangles(h.data_ptr() if hands else None, f.data_ptr() if faces else None,
                                                          keypoints.data_ptr(), count.data_ptr(), n, cap, int(pose_model),
                                                          _stream()))
    return h, f


class HandDetector:
    def __init__(self, poseModel: int):
        self.mPoseModel = int(poseModel)

    def detectHands(self, keypoints: torch.Tensor, count: torch.Tensor) -> torch.Tensor:
        return hand_face_rectangles_gpu(keypoints, count, self.mPoseModel, True, False)[0]


class FaceDetector:
    def __init__(self, poseModel: int):
        self.mPoseModel = int(poseModel)

    def detectFaces(self, keypoints: torch.Tensor, count: torch.Tensor) -> torch.Tensor:
        return hand_face_rectangles_gpu(keypoints, count, self.mPoseModel, False, True)[1]
