"""Output serialisation of the sharded stream (libopb200: csrc/host/people_json.cpp; host only).

Mirrors (reference files relative to zengjianping/openpose; same names and argument meaning):
  savePeopleJson        <- src/openpose/filestream/fileStream.cpp:289-344 (+ addKeypointsToJson :20-78,
                           addCandidatesToJson :80-128, JsonOfstream src/openpose/filestream/jsonOfstream.cpp)
  PeopleJsonSaver.save  <- src/openpose/filestream/peopleJsonSaver.cpp:15-30 (FileSaver: fileSaver.cpp:6-48)
  CocoJsonSaver         <- src/openpose/filestream/cocoJsonSaver.cpp:28-277
  peopleJsonBatch       <- WPeopleJsonSaver::workConsumer, include/openpose/filestream/wPeopleJsonSaver.hpp:55-100,
                           for every frame of a connector batch at once
keypointVector: list of (array, name): array = None / empty (an empty op::Array), 1-D (people) or 3-D
(people, parts, dims) float32. candidates: list (one entry per body part) of (n, 3) arrays, or None / [].
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence, Tuple

import numpy as np

from . import _lib
from ._lib import Opb200Error


class _JsonArray(C.Structure):
    _fields_ = [("name", C.c_char_p), ("data", C.c_void_p), ("people", C.c_int), ("per_person", C.c_int)]


_lib.register({
    "opb200_people_json": (C.c_int, [C.POINTER(_JsonArray), C.c_int, C.c_void_p, C.POINTER(C.c_int), C.c_int, C.c_int,
                                     C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]),
    "opb200_people_json_save": (C.c_int, [C.c_char_p, C.POINTER(_JsonArray), C.c_int, C.c_void_p, C.POINTER(C.c_int),
                                          C.c_int, C.c_int]),
    "opb200_people_json_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p,
                                           C.c_size_t, C.POINTER(C.c_size_t)]),
    "opb200_coco_json_open": (C.c_int, [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                        C.POINTER(C.c_void_p)]),
    "opb200_coco_json_record": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_char_p,
                                          C.c_ulonglong]),
    "opb200_coco_json_close": (C.c_int, [C.c_void_p]),
})


def _pack(keypointVector, candidates):
    keep = []
    arr = (_JsonArray * max(1, len(keypointVector)))()
    for i, (a, name) in enumerate(keypointVector):
        if a is None or np.size(a) == 0:
            arr[i] = _JsonArray(name.encode(), None, 0, 0)
            continue
        a = np.ascontiguousarray(a, np.float32)
        if a.ndim not in (1, 3):   # fileStream.cpp:314-317
            raise Opb200Error("keypointVector.getNumberDimensions() != 1 && != 3.")
        keep.append(a)
        per = 1 if a.ndim == 1 else a.shape[1] * a.shape[2]
        arr[i] = _JsonArray(name.encode(), a.ctypes.data, a.shape[0], per)
    n_parts = len(candidates) if candidates is not None else 0
    counts = (C.c_int * max(1, n_parts))()
    flat = np.zeros((0, 3), np.float32)
    if n_parts:
        parts = [np.asarray(c, np.float32).reshape(-1, 3) for c in candidates]
        for i, c in enumerate(parts):
            counts[i] = c.shape[0]
        flat = np.ascontiguousarray(np.concatenate(parts, 0), np.float32)
    keep.append(flat)
    return arr, len(keypointVector), flat, counts, n_parts, keep


def peopleJsonBytes(keypointVector, candidates=None, humanReadable: bool = True) -> bytes:
    """The bytes savePeopleJson writes to its std::ofstream."""
    arr, n, flat, counts, n_parts, _keep = _pack(keypointVector, candidates)
    L = _lib.lib()
    length = C.c_size_t(0)
    _lib.check(L.opb200_people_json(arr, n, flat.ctypes.data, counts, n_parts, int(humanReadable), None, 0,
                                    C.byref(length)))
    buf = C.create_string_buffer(max(1, length.value))
    _lib.check(L.opb200_people_json(arr, n, flat.ctypes.data, counts, n_parts, int(humanReadable), buf, length.value,
                                    C.byref(length)))
    return buf.raw[:length.value]


def savePeopleJson(keypointVector, candidates, fileName: str, humanReadable: bool) -> None:
    arr, n, flat, counts, n_parts, _keep = _pack(keypointVector, candidates)
    _lib.check(_lib.lib().opb200_people_json_save(fileName.encode(), arr, n, flat.ctypes.data, counts, n_parts,
                                                  int(humanReadable)))


class PeopleJsonSaver:
    def __init__(self, directoryPath: str):
        d = directoryPath.replace("\\", "/")          # formatAsDirectory, utilities/fileSystem.cpp
        if d and not d.endswith("/"):
            d += "/"
        self.mDirectoryPath = d
        if d:
            os.makedirs(d, exist_ok=True)              # makeDirectory

    def save(self, keypointVector, candidates, fileName: str, humanReadable: bool = True) -> None:
        savePeopleJson(keypointVector, candidates, self.mDirectoryPath + fileName + ".json", humanReadable)


def peopleJsonBatch(keypoints: np.ndarray, counts: np.ndarray, humanReadable: bool = False) -> List[bytes]:
    """keypoints (frames, capacity, parts, 3) + counts (frames) on the HOST (the connector's output after its D2H
    copy) -> one JSON document per frame, what WPeopleJsonSaver writes for a body-only Datum."""
    kp = np.ascontiguousarray(keypoints, np.float32)
    cnt = np.ascontiguousarray(counts, np.int32)
    if kp.ndim != 4 or kp.shape[3] != 3 or cnt.shape != (kp.shape[0],):
        raise Opb200Error("peopleJsonBatch: keypoints must be (frames, capacity, parts, 3), counts (frames)")
    frames, cap, parts = kp.shape[:3]
    L = _lib.lib()
    offs = (C.c_size_t * (frames + 1))()
    _lib.check(L.opb200_people_json_batch(kp.ctypes.data, cnt.ctypes.data, frames, cap, parts, int(humanReadable),
                                          None, 0, offs))
    total = offs[frames]
    buf = C.create_string_buffer(max(1, total))
    _lib.check(L.opb200_people_json_batch(kp.ctypes.data, cnt.ctypes.data, frames, cap, parts, int(humanReadable),
                                          buf, total, offs))
    raw = buf.raw
    return [raw[offs[f]:offs[f + 1]] for f in range(frames)]


class CocoJsonSaver:
    """cocoJsonFormat: op::CocoJsonFormat (0 body, 4 foot, 5 car); poseModel: op::PoseModel (0 BODY_25, 1 COCO_18)."""

    def __init__(self, filePathToSave: str, poseModel: int = 0, humanReadable: bool = True, cocoJsonVariants: int = 1,
                 cocoJsonFormat: int = 0, cocoJsonVariant: int = 0):
        self._h = C.c_void_p()
        _lib.check(_lib.lib().opb200_coco_json_open(filePathToSave.encode(), int(poseModel), int(humanReadable),
                                                    int(cocoJsonVariants), int(cocoJsonFormat), int(cocoJsonVariant),
                                                    C.byref(self._h)))

    def record(self, poseKeypoints, poseScores, imageName: str, frameNumber: int = 0) -> None:
        kp = np.ascontiguousarray(poseKeypoints, np.float32)
        sc = np.ascontiguousarray(poseScores, np.float32)
        people = kp.shape[0] if kp.size else 0
        if people != sc.size:   # cocoJsonSaver.cpp:101-102
            raise Opb200Error("Dimension mismatch between poseKeypoints and poseScores.")
        parts = kp.shape[1] if kp.size else 0
        _lib.check(_lib.lib().opb200_coco_json_record(self._h, kp.ctypes.data, sc.ctypes.data, people, parts,
                                                      imageName.encode(), int(frameNumber)))

    def close(self) -> None:
        if self._h:
            h, self._h = self._h, C.c_void_p()
            _lib.check(_lib.lib().opb200_coco_json_close(h))

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
