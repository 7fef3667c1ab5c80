"""Host-side mirror of the reference's body-pose extractor for the hot path: op::PoseExtractorCaffe /
op::PoseExtractorNet driving net -> resizeAndMerge -> nms -> bodyPartConnector, with the reshape rules of
the four Caffe-style layer classes and the heat-map / candidate getters.

Mirrors (reference files relative to zengjianping/openpose, same names / argument meaning / error text):
  PoseExtractorB200.forwardPass        <- PoseExtractorCaffe::forwardPass      src/openpose/pose/poseExtractorCaffe.cpp:200-345
  reshape rules                        <- reshapePoseExtractorCaffe            poseExtractorCaffe.cpp:35-71
      resize_and_merge_top_shape       <- ResizeAndMergeCaffe::Reshape         src/openpose/net/resizeAndMergeCaffe.cpp:58-110
      nms_top_shape                    <- NmsCaffe::Reshape                    src/openpose/net/nmsCaffe.cpp:105-154
      maximum_top_shape                <- MaximumCaffe::Reshape                src/openpose/net/maximumCaffe.cpp:60-75
  getHeatMapsCopy / getCandidatesCopy  <- PoseExtractorNet                     src/openpose/pose/poseExtractorNet.cpp:106-282
  get / set / getScaleNetToOutput ...  <- PoseExtractorNet                     poseExtractorNet.cpp:284-360
  resize_get_scale_factor              <- resizeGetScaleFactor                 src/openpose/utilities/openCv.cpp:182-195
Everything numeric runs on the device through the C ABI (openpose_b200.net / openpose_b200.post); this file is
the orchestration only (shapes, scale factors, property plumbing), as in the reference.
"""
from __future__ import annotations

import math
from enum import IntEnum
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import net as netmod
from . import post
from ._lib import Opb200Error, POSE_BODY_25, POSE_COCO_18, SEMANTICS_CPU

POSE_MAX_PEOPLE = post.POSE_MAX_PEOPLE


class HeatMapType(IntEnum):          # include/openpose/core/enumClasses.hpp
    Parts = 0
    Background = 1
    PAFs = 2


class ScaleMode(IntEnum):            # include/openpose/core/enumClasses.hpp
    InputResolution = 0
    NetOutputResolution = 1
    OutputResolution = 2
    ZeroToOne = 3
    ZeroToOneFixedAspect = 4
    PlusMinusOne = 5
    PlusMinusOneFixedAspect = 6
    UnsignedChar = 7
    NoScale = 8


class PoseProperty(IntEnum):         # include/openpose/pose/enumClasses.hpp:66-77
    NMSThreshold = 0
    ConnectInterMinAboveThreshold = 1
    ConnectInterThreshold = 2
    ConnectMinSubsetCnt = 3
    ConnectMinSubsetScore = 4


def positive_int_round(a: float) -> int:
    """positiveIntRound (include/openpose/utilities/fastMath.hpp): int(a + 0.5f)."""
    return int(a + 0.5)


def resize_get_scale_factor(initial_size: Tuple[int, int], target_size: Tuple[int, int]) -> float:
    """(x, y) sizes -> min((tw-1)/(iw-1), (th-1)/(ih-1)) in double precision."""
    rw = (target_size[0] - 1) / float(initial_size[0] - 1)
    rh = (target_size[1] - 1) / float(initial_size[1] - 1)
    return min(rw, rh)


def pose_net_decrease_factor(pose_model: int) -> float:
    """getPoseNetDecreaseFactor (poseParameters.cpp:630-641): 8 for every shipped body model."""
    return 8.0


def pose_number_body_parts(pose_model: int) -> int:
    return post.model_dims(pose_model)[0]


def resize_and_merge_top_shape(bottom_shape: Sequence[int], net_factor: float, scale_factor: float,
                               merge_first_dimension: bool) -> List[int]:
    """H' = round((H * netFactor - 1) * scaleFactor) + 1 in float arithmetic (the -1/+1 is the 0-based index rule)."""
    top = list(bottom_shape)
    top[0] = 1 if merge_first_dimension else bottom_shape[0]
    f = np.float32
    for d in (2, 3):
        v = (f(top[d]) * f(net_factor) - f(1.0)) * f(scale_factor)
        top[d] = int(math.floor(float(v) + 0.5) if v >= 0 else -math.floor(-float(v) + 0.5)) + 1   # std::round
    return top


def nms_top_shape(bottom_shape: Sequence[int], max_peaks: int, output_channels: int) -> List[int]:
    """(N, outputChannels (<= C, or C-1 when negative), maxPeaks + 1, 3)."""
    channels = output_channels if output_channels > 0 else bottom_shape[1] - 1
    return [bottom_shape[0], channels, max_peaks + 1, 3]


def maximum_top_shape(bottom_shape: Sequence[int]) -> List[int]:
    """(N, 1, C - 1, 3): the background channel is dropped."""
    return [bottom_shape[0], 1, bottom_shape[1] - 1, 3]


class PoseExtractorB200:
    """Body-pose extractor on one GPU. One instance per GPU, initialised and used on one thread, as the reference's."""

    def __init__(self, poseModel: int, caffeProto: str, caffeTrainedModel: str, gpuId: int = 0,
                 heatMapTypes: Sequence[HeatMapType] = (), heatMapScaleMode: ScaleMode = ScaleMode.ZeroToOneFixedAspect,
                 addPartCandidates: bool = False, maximizePositives: bool = False, enableNet: bool = True,
                 upsamplingRatio: float = 0.0, semantics: int = SEMANTICS_CPU, capacity: int = 512):
        if poseModel not in (POSE_BODY_25, POSE_COCO_18):
            raise Opb200Error("Model not implemented for the device connector (BODY_25 and COCO_18 only).")
        self.mPoseModel, self.mGpuId = poseModel, gpuId
        self.mProto, self.mModel = caffeProto, caffeTrainedModel
        self.mHeatMapTypes, self.mHeatMapScaleMode = list(heatMapTypes), ScaleMode(heatMapScaleMode)
        self.mAddPartCandidates, self.mMaximizePositives = addPartCandidates, maximizePositives
        self.mEnableNet, self.mUpsamplingRatio = enableNet, float(upsamplingRatio)
        self.mSemantics, self.mCapacity = semantics, capacity
        d = post.pose_defaults(poseModel, maximizePositives)     # poseParameters.cpp:677-756
        self.mProperties = {PoseProperty.NMSThreshold: d["nms_threshold"],
                            PoseProperty.ConnectInterMinAboveThreshold: d["inter_min_above"],
                            PoseProperty.ConnectInterThreshold: d["inter_threshold"],
                            PoseProperty.ConnectMinSubsetCnt: d["min_subset_cnt"],
                            PoseProperty.ConnectMinSubsetScore: d["min_subset_score"]}
        self.spNets: List[netmod.NetB200] = []
        self.mNetInput4DSizes: List[Tuple[int, ...]] = []
        self.mNetOutputSize = (0, 0)
        self.mScaleNetToOutput = 1.0
        self.spHeatMapsBlob: Optional[torch.Tensor] = None
        self.spPeaksBlob: Optional[torch.Tensor] = None
        self.mPoseKeypoints = np.zeros((0, pose_number_body_parts(poseModel), 3), np.float32)
        self.mPoseScores = np.zeros((0,), np.float32)
        self._initialized = False

    # ---- PoseExtractorNet plumbing ------------------------------------------------------------------
    def initializationOnThread(self) -> None:
        if self.mEnableNet:
            self._add_net()
        self._initialized = True

    def _add_net(self) -> None:          # addCaffeNetOnThread (poseExtractorCaffe.cpp:73-98)
        n = netmod.NetB200(self.mProto, self.mModel, gpuId=self.mGpuId)
        n.initializationOnThread()
        self.spNets.append(n)

    def get(self, prop: PoseProperty) -> float:
        return self.mProperties[PoseProperty(prop)]

    def set(self, prop: PoseProperty, value: float) -> None:
        self.mProperties[PoseProperty(prop)] = value

    def increase(self, prop: PoseProperty, value: float) -> None:
        self.mProperties[PoseProperty(prop)] += value

    def getScaleNetToOutput(self) -> float:
        return self.mScaleNetToOutput

    def getPoseKeypoints(self) -> np.ndarray:
        return self.mPoseKeypoints

    def getPoseScores(self) -> np.ndarray:
        return self.mPoseScores

    def getHeatMapSize(self) -> List[int]:
        return list(self.spHeatMapsBlob.shape) if self.spHeatMapsBlob is not None else []

    def getHeatMapGpuConstPtr(self) -> torch.Tensor:
        return self.spHeatMapsBlob

    def getCandidatesGpuConstPtr(self) -> torch.Tensor:
        return self.spPeaksBlob

    # ---- forward ------------------------------------------------------------------------------------
    def forwardPass(self, inputNetData: Sequence, inputDataSize: Tuple[int, int],
                    scaleInputToNetInputs: Sequence[float], poseNetOutput=None) -> None:
        """inputNetData: one (1,3,H_i,W_i) float32 array per scale (numpy, CPU or CUDA tensor); inputDataSize =
        (width, height) of the original frame; poseNetOutput: (1,C,h,w) custom heat-maps when enableNet is False."""
        if not self._initialized:
            raise Opb200Error("PoseExtractorB200: initializationOnThread() has not been called")
        empty = lambda a: a is None or (hasattr(a, "numel") and a.numel() == 0) or getattr(a, "size", 1) == 0  # noqa: E731
        if len(inputNetData) == 0 or any(empty(a) for a in inputNetData):
            raise Opb200Error("Empty inputNetData.")
        if len(inputNetData) != len(scaleInputToNetInputs):
            raise Opb200Error("Size(inputNetData) must be same than size(scaleInputToNetInputs).")
        if empty(poseNetOutput) != self.mEnableNet:
            msg = (". Either use OpenPose default network (`--body 1`) or fill the `poseNetOutput` argument (only 1 "
                   "of those 2, not both).")
            if empty(poseNetOutput):
                raise Opb200Error("The argument poseNetOutput cannot be empty if mEnableNet is true" + msg)
            raise Opb200Error("The argument poseNetOutput is not empty and you have also explicitly chosen to run "
                              "the OpenPose network" + msg)
        dev = torch.device("cuda", self.mGpuId)
        number_scales = len(inputNetData)
        # 1. deep net, one Net per scale (poseExtractorCaffe.cpp:240-249)
        if self.mEnableNet:
            while len(self.spNets) < number_scales:
                self._add_net()
            outs = []
            for i, x in enumerate(inputNetData):
                self.spNets[i].forwardPass(x)
                outs.append(self.spNets[i].getOutputBlobArray())
        else:
            if number_scales != 1:
                raise Opb200Error("Size(inputNetData) must match the provided heatmaps batch size (%d vs. 1)."
                                  % number_scales)
            outs = [torch.as_tensor(np.ascontiguousarray(poseNetOutput, np.float32) if not
                                    isinstance(poseNetOutput, torch.Tensor) else poseNetOutput).to(dev)]
        # reshape blobs on size change (poseExtractorCaffe.cpp:262-288)
        sizes = [tuple(int(v) for v in a.shape) for a in inputNetData]
        parts = pose_number_body_parts(self.mPoseModel)
        if sizes != self.mNetInput4DSizes or self.spHeatMapsBlob is None:
            self.mNetInput4DSizes = sizes
            net_factor = (pose_net_decrease_factor(self.mPoseModel) if self.mUpsamplingRatio <= 0.0
                          else self.mUpsamplingRatio)
            top = resize_and_merge_top_shape(list(outs[0].shape), net_factor, 1.0 / 1.0, True)
            self.spHeatMapsBlob = torch.empty(top, dtype=torch.float32, device=dev)
            self.spPeaksBlob = torch.zeros(nms_top_shape(top, POSE_MAX_PEOPLE, parts), dtype=torch.float32, device=dev)
            ratio = 1.0 if self.mUpsamplingRatio <= 0.0 else self.mUpsamplingRatio / pose_net_decrease_factor(
                self.mPoseModel)
            self.mNetOutputSize = (positive_int_round(ratio * sizes[0][3]), positive_int_round(ratio * sizes[0][2]))
        # 2. resize heat maps + merge scales; ratios are converted double -> float as in the reference
        ratios = [float(np.float32(r)) for r in scaleInputToNetInputs]
        post.resize_and_merge_gpu(self.spHeatMapsBlob, outs, ratios)
        scale_producer_to_net = resize_get_scale_factor(inputDataSize, self.mNetOutputSize)
        net_size = (positive_int_round(scale_producer_to_net * inputDataSize[0]),
                    positive_int_round(scale_producer_to_net * inputDataSize[1]))
        self.mScaleNetToOutput = float(np.float32(resize_get_scale_factor(net_size, inputDataSize)))
        # 3. peaks by non-maximum suppression
        nms_threshold = float(np.float32(self.get(PoseProperty.NMSThreshold)))
        nms_offset = float(np.float32(0.5 / float(self.mScaleNetToOutput)))
        post.nms_gpu(self.spPeaksBlob, self.spHeatMapsBlob, nms_threshold, (nms_offset, nms_offset), self.mSemantics)
        # 4. connecting body parts
        params = post.make_connect_params(
            self.mPoseModel, scale_factor=self.mScaleNetToOutput, maximize_positives=self.mMaximizePositives,
            semantics=self.mSemantics,
            overrides=dict(nms_threshold=self.get(PoseProperty.NMSThreshold),
                           inter_min_above=self.get(PoseProperty.ConnectInterMinAboveThreshold),
                           inter_threshold=self.get(PoseProperty.ConnectInterThreshold),
                           min_subset_cnt=int(self.get(PoseProperty.ConnectMinSubsetCnt)),
                           min_subset_score=self.get(PoseProperty.ConnectMinSubsetScore)))
        kp, sc, cnt = post.connect_body_parts_gpu(self.spHeatMapsBlob, self.spPeaksBlob, params,
                                                  capacity=self.mCapacity)
        n = int(cnt[0].item())
        self.mPoseKeypoints = kp[0, :n].cpu().numpy()
        self.mPoseScores = sc[0, :n].cpu().numpy()

    # ---- getters (poseExtractorNet.cpp:106-282) --------------------------------------------------------
    def getHeatMapsCopy(self) -> np.ndarray:
        """(channels, H, W) host copy of the selected channel groups, range-remapped per the ScaleMode."""
        if not self.mHeatMapTypes or self.spHeatMapsBlob is None:
            return np.zeros((0,), np.float32)
        parts = pose_number_body_parts(self.mPoseModel)
        pairs = post.model_dims(self.mPoseModel)[1]
        mask = sum(1 << int(t) for t in set(self.mHeatMapTypes))
        hm = self.spHeatMapsBlob[0:1]
        with torch.cuda.device(hm.device):
            return post.heatmaps_export_gpu(hm, parts, True, 2 * pairs, mask, self.mHeatMapScaleMode)[0].numpy()

    def getCandidatesCopy(self) -> List[List[Tuple[float, float, float]]]:
        """Per body part: [(x * scaleNetToOutput, y * scaleNetToOutput, score)] of the first count peaks."""
        if not self.mAddPartCandidates or self.spPeaksBlob is None:
            return []
        pk = self.spPeaksBlob[0].cpu().numpy()
        s = np.float32(self.mScaleNetToOutput)
        out = []
        for part in range(pk.shape[0]):
            n = int(round(float(pk[part, 0, 0])))
            rows = pk[part, 1:1 + n]
            out.append([(float(r[0] * s), float(r[1] * s), float(r[2])) for r in rows])
        return out
