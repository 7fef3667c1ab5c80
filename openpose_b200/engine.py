"""Python front-end of the multi-GPU body-pose engine (C ABI: include/opb200.h, opb200_engine_*): u8 BGR frames in,
people out, frame order kept. It mirrors what the reference's Wrapper does for this path -- one PoseExtractor per GPU on
its own thread, frames dealt round-robin, results re-ordered by frame id (include/openpose/wrapper/wrapperAuxiliary.hpp:
181-219,326-338, include/openpose/thread/wQueueOrderer.hpp:12-80) -- with the threads, streams and staging inside
libopb200.so. numpy only: the engine owns all device memory."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import numpy as np

from . import _lib
from ._lib import ConnectParams, Opb200Error, SEMANTICS_CPU


class EngineConfig(C.Structure):
    _fields_ = [("pose_model", C.c_int), ("net_w", C.c_int), ("net_h", C.c_int), ("frame_w", C.c_int),
                ("frame_h", C.c_int), ("batch", C.c_int), ("lanes", C.c_int), ("capacity", C.c_int),
                ("scale_input_to_net", C.c_float), ("nms_threshold", C.c_float), ("connect", ConnectParams)]


class CropEngineConfig(C.Structure):
    _fields_ = [("heat_channels", C.c_int), ("net_side", C.c_int), ("frame_w", C.c_int), ("frame_h", C.c_int),
                ("batch", C.c_int), ("lanes", C.c_int)]


_lib.register({
    "opb200_crop_engine_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_char_p, C.c_char_p, C.POINTER(C.c_int), C.c_int,
                                            C.POINTER(CropEngineConfig)]),
    "opb200_crop_engine_matrix_offset": (C.c_size_t, [C.c_void_p]),
    "opb200_crop_engine_submit": (C.c_int, [C.c_void_p, C.c_ulonglong, C.c_void_p, C.c_void_p, C.c_int]),
    "opb200_crop_engine_next": (C.c_int, [C.c_void_p, C.POINTER(C.c_ulonglong), C.POINTER(C.c_int), C.c_void_p]),
    "opb200_engine_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_char_p, C.c_char_p, C.POINTER(C.c_int), C.c_int,
                                       C.POINTER(EngineConfig)]),
    "opb200_engine_destroy": (None, [C.c_void_p]),
    "opb200_engine_acquire": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "opb200_engine_commit": (C.c_int, [C.c_void_p, C.c_ulonglong, C.c_int]),
    "opb200_engine_submit": (C.c_int, [C.c_void_p, C.c_ulonglong, C.c_void_p, C.c_int]),
    "opb200_engine_next": (C.c_int, [C.c_void_p, C.POINTER(C.c_ulonglong), C.POINTER(C.c_int), C.c_void_p, C.c_void_p,
                                     C.c_void_p]),
    "opb200_engine_in_flight": (C.c_int, [C.c_void_p]),
})


class PoseEngine:
    """frames (n, frame_h, frame_w, 3) u8 BGR -> keypoints (n, capacity, parts, 3), scores (n, capacity), counts (n)."""

    def __init__(self, prototxt: str, caffemodel: str, gpus: Sequence[int] = (0,), pose_model: int = 0,
                 net_size: Tuple[int, int] = (656, 368), frame_size: Tuple[int, int] = (1280, 720), batch: int = 44,
                 lanes: int = 2, capacity: int = 256, semantics: int = SEMANTICS_CPU, maximize_positives: bool = False,
                 overrides: Optional[dict] = None):
        from .op_input import ScaleAndSizeExtractor
        from .post import make_connect_params, pose_defaults
        from .pose_extractor import positive_int_round, resize_get_scale_factor
        scales, sizes, _, _ = ScaleAndSizeExtractor(net_size).extract(frame_size)
        assert tuple(sizes[0]) == tuple(net_size), (sizes, net_size)
        # scaleNetToOutput as PoseExtractorCaffe::forwardPass derives it (poseExtractorCaffe.cpp:306-310)
        s = resize_get_scale_factor(frame_size, net_size)
        net_scaled = (positive_int_round(s * frame_size[0]), positive_int_round(s * frame_size[1]))
        self.scale_net_to_output = float(np.float32(resize_get_scale_factor(net_scaled, frame_size)))
        d = pose_defaults(pose_model, maximize_positives)
        if overrides:
            d.update(overrides)
        self.cfg = EngineConfig(pose_model, net_size[0], net_size[1], frame_size[0], frame_size[1], batch, lanes, capacity,
                                float(scales[0]), float(d["nms_threshold"]),
                                make_connect_params(pose_model, self.scale_net_to_output, maximize_positives, semantics,
                                                    overrides=overrides))
        self.parts = 25 if pose_model == 0 else 18
        self.batch, self.capacity = batch, capacity
        self.frame_shape = (frame_size[1], frame_size[0], 3)
        L = _lib.lib()
        ids = (C.c_int * len(gpus))(*[int(g) for g in gpus])
        self._h = C.c_void_p()
        _lib.check(L.opb200_engine_create(C.byref(self._h), prototxt.encode(), caffemodel.encode(), ids, len(gpus),
                                          C.byref(self.cfg)))
        self._kp = np.empty((batch, capacity, self.parts, 3), np.float32)
        self._sc = np.empty((batch, capacity), np.float32)
        self._cnt = np.empty((batch,), np.int32)

    def acquire(self) -> np.ndarray:
        """The next pinned staging buffer as a (batch, frame_h, frame_w, 3) u8 array to write frames into."""
        p = C.c_void_p()
        _lib.check(_lib.lib().opb200_engine_acquire(self._h, C.byref(p)))
        n = self.batch * int(np.prod(self.frame_shape))
        buf = (C.c_ubyte * n).from_address(p.value)
        return np.frombuffer(buf, np.uint8).reshape((self.batch,) + self.frame_shape)

    def commit(self, first_frame_id: int, n_frames: int) -> None:
        _lib.check(_lib.lib().opb200_engine_commit(self._h, int(first_frame_id), int(n_frames)))

    def submit(self, first_frame_id: int, frames: np.ndarray) -> None:
        frames = np.ascontiguousarray(frames, np.uint8)
        if frames.ndim != 4 or frames.shape[1:] != self.frame_shape:
            raise Opb200Error("frames must be (n, %d, %d, 3) uint8" % self.frame_shape[:2])
        _lib.check(_lib.lib().opb200_engine_submit(self._h, int(first_frame_id), frames.ctypes.data, frames.shape[0]))

    def next(self, copy: bool = True):
        """(first_frame_id, keypoints (n,capacity,parts,3), scores (n,capacity), counts (n)) of the oldest batch."""
        fid, n = C.c_ulonglong(), C.c_int()
        _lib.check(_lib.lib().opb200_engine_next(self._h, C.byref(fid), C.byref(n), self._kp.ctypes.data,
                                                 self._sc.ctypes.data, self._cnt.ctypes.data))
        k = n.value
        out = (self._kp[:k], self._sc[:k], self._cnt[:k])
        if copy:
            out = tuple(a.copy() for a in out)
        return (int(fid.value),) + out

    def in_flight(self) -> int:
        return int(_lib.lib().opb200_engine_in_flight(self._h))

    def close(self) -> None:
        if self._h:
            _lib.lib().opb200_engine_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class CropEngine:
    """Hand / face crops over one or more GPUs (C ABI: opb200_crop_engine_*): one u8 BGR frame + the 2x3 inverse affine
    maps of its crops in, (n_crops, parts, 3) maxima in crop coordinates out, submission order kept. The device part of
    HandExtractorCaffe::forwardPass / FaceExtractorCaffe::forwardPass (src/openpose/hand/handExtractorCaffe.cpp:305-464,
    src/openpose/face/faceExtractorCaffe.cpp:175-300) for all crops of a batch at once."""

    def __init__(self, prototxt: str, caffemodel: str, heat_channels: int, gpus: Sequence[int] = (0,),
                 frame_size: Tuple[int, int] = (1280, 720), net_side: int = 368, batch: int = 16, lanes: int = 2):
        self.cfg = CropEngineConfig(int(heat_channels), int(net_side), frame_size[0], frame_size[1], int(batch), int(lanes))
        self.parts, self.batch = int(heat_channels) - 1, int(batch)
        self.frame_shape = (frame_size[1], frame_size[0], 3)
        L = _lib.lib()
        ids = (C.c_int * len(gpus))(*[int(g) for g in gpus])
        self._h = C.c_void_p()
        _lib.check(L.opb200_crop_engine_create(C.byref(self._h), prototxt.encode(), caffemodel.encode(), ids, len(gpus),
                                               C.byref(self.cfg)))
        self._peaks = np.empty((batch, self.parts, 3), np.float32)

    def submit(self, frame_id: int, frame: np.ndarray, matrices: np.ndarray) -> None:
        frame = np.ascontiguousarray(frame, np.uint8)
        matrices = np.ascontiguousarray(matrices, np.float64).reshape(-1, 2, 3)
        if frame.shape != self.frame_shape:
            raise Opb200Error("frame must be (%d, %d, 3) uint8" % self.frame_shape[:2])
        _lib.check(_lib.lib().opb200_crop_engine_submit(self._h, int(frame_id), frame.ctypes.data, matrices.ctypes.data,
                                                        matrices.shape[0]))

    def next(self):
        """(frame_id, peaks (n_crops, parts, 3)) of the oldest submitted batch."""
        fid, n = C.c_ulonglong(), C.c_int()
        _lib.check(_lib.lib().opb200_crop_engine_next(self._h, C.byref(fid), C.byref(n), self._peaks.ctypes.data))
        return int(fid.value), self._peaks[:n.value].copy()

    def in_flight(self) -> int:
        return int(_lib.lib().opb200_engine_in_flight(self._h))

    def close(self) -> None:
        if self._h:
            _lib.lib().opb200_engine_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
