"""Synthetic inputs for tests and benchmarks (there is no network for datasets or trained weights).

* planted_maps(): heat-maps + PAFs with planted people: every body
  part a Gaussian blob, every limb a band of unit vectors, plus noise. Channel layout follows the
  reference's net_output: [0..parts-1] part maps, [parts] background, then 2 PAF channels per limb
  at parts+1+mapIdx (reference: src/openpose/net/bodyPartConnectorBase.cpp:174,299-302).
* frames(): u8 BGR frames and their network-input tensors (x/256-0.5, CHW), as
  src/openpose/core/cvMatToOpInput.cpp:63-159 produces them (resize itself is out of scope: frames are
  generated at network resolution).
numpy only; nothing here is on the timed path.
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np

BODY_25, COCO_18 = 0, 1

# (part pairs, PAF map index) per model; data from src/openpose/pose/poseParameters.cpp:253-262,416-424
_PAIRS = {
    BODY_25: [1, 8, 1, 2, 1, 5, 2, 3, 3, 4, 5, 6, 6, 7, 8, 9, 9, 10, 10, 11, 8, 12, 12, 13, 13, 14, 1, 0, 0, 15, 15, 17,
              0, 16, 16, 18, 2, 17, 5, 18, 14, 19, 19, 20, 14, 21, 11, 22, 22, 23, 11, 24],
    COCO_18: [1, 2, 1, 5, 2, 3, 3, 4, 5, 6, 6, 7, 1, 8, 8, 9, 9, 10, 1, 11, 11, 12, 12, 13, 1, 0, 0, 14, 14, 16, 0, 15,
              15, 17, 2, 16, 5, 17],
}
_MAPIDX = {
    BODY_25: [0, 1, 14, 15, 22, 23, 16, 17, 18, 19, 24, 25, 26, 27, 6, 7, 2, 3, 4, 5, 8, 9, 10, 11, 12, 13, 30, 31, 32,
              33, 36, 37, 34, 35, 38, 39, 20, 21, 28, 29, 40, 41, 42, 43, 44, 45, 46, 47, 48, 49, 50, 51],
    COCO_18: [12, 13, 20, 21, 14, 15, 16, 17, 22, 23, 24, 25, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 28, 29, 30, 31, 34,
              35, 32, 33, 36, 37, 18, 19, 26, 27],
}
_PARTS = {BODY_25: 25, COCO_18: 18}


def model_tables(model: int) -> Tuple[int, List[int], List[int]]:
    return _PARTS[model], _PAIRS[model], _MAPIDX[model]


def net_output_channels(model: int) -> int:
    parts, pairs, _ = model_tables(model)
    return parts + 1 + len(pairs)


def plant_people(rng: np.random.Generator, model: int, n_people: int, h: int, w: int,
                 limb_len=(18.0, 48.0), drop_prob=0.0) -> np.ndarray:
    """(n_people, parts, 3) = (x, y, amplitude); amplitude 0 = part absent. Coordinates in map pixels."""
    parts, pairs, _ = model_tables(model)
    out = np.zeros((n_people, parts, 3), np.float32)
    cols = max(1, int(np.ceil(np.sqrt(n_people * w / max(h, 1)))))
    rows = max(1, int(np.ceil(n_people / cols)))
    scale = min(1.0, min(w / cols, h / rows) / (2.5 * limb_len[1]))
    for p in range(n_people):
        cx = (p % cols + 0.5) * w / cols + rng.uniform(-0.15, 0.15) * w / cols
        cy = (p // cols + 0.5) * h / rows + rng.uniform(-0.15, 0.15) * h / rows
        pos = {1: (cx, cy)}
        pending = list(range(len(pairs) // 2))
        while pending:
            progressed = False
            for l in list(pending):
                a, b = pairs[2 * l], pairs[2 * l + 1]
                src, dst = (a, b) if a in pos else ((b, a) if b in pos else (None, None))
                if src is None:
                    continue
                pending.remove(l)
                progressed = True
                if dst in pos:
                    continue
                ang = rng.uniform(0, 2 * np.pi)
                ln = rng.uniform(*limb_len) * scale
                x = float(np.clip(pos[src][0] + ln * np.cos(ang), 4, w - 5))
                y = float(np.clip(pos[src][1] + ln * np.sin(ang), 4, h - 5))
                pos[dst] = (x, y)
            if not progressed:
                break
        for part, (x, y) in pos.items():
            if drop_prob > 0 and rng.uniform() < drop_prob:
                continue
            out[p, part] = (x, y, rng.uniform(0.3, 1.0))
    return out


def planted_maps(model: int, h: int, w: int, n_people: int, seed: int, sigma: float = 7.0,
                 paf_halfwidth: float = 4.0, noise: float = 0.01, drop_prob: float = 0.0,
                 limb_len=(18.0, 48.0)) -> Tuple[np.ndarray, np.ndarray]:
    """Returns (maps (C, h, w) float32, people (n_people, parts, 3))."""
    rng = np.random.default_rng(seed)
    parts, pairs, mapidx = model_tables(model)
    channels = net_output_channels(model)
    people = plant_people(rng, model, n_people, h, w, limb_len=limb_len, drop_prob=drop_prob)
    maps = np.zeros((channels, h, w), np.float32)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    r = int(np.ceil(4 * sigma))
    for p in range(n_people):
        for part in range(parts):
            x, y, a = people[p, part]
            if a <= 0:
                continue
            x0, x1 = max(0, int(x) - r), min(w, int(x) + r + 1)
            y0, y1 = max(0, int(y) - r), min(h, int(y) + r + 1)
            g = a * np.exp(-((xx[y0:y1, x0:x1] - x) ** 2 + (yy[y0:y1, x0:x1] - y) ** 2) / (2 * sigma * sigma))
            np.maximum(maps[part, y0:y1, x0:x1], g, out=maps[part, y0:y1, x0:x1])
        for l in range(len(pairs) // 2):
            a, b = pairs[2 * l], pairs[2 * l + 1]
            if people[p, a, 2] <= 0 or people[p, b, 2] <= 0:
                continue
            ax, ay = people[p, a, :2]
            bx, by = people[p, b, :2]
            dx, dy = bx - ax, by - ay
            ln = float(np.hypot(dx, dy))
            if ln < 1e-3:
                continue
            ux, uy = dx / ln, dy / ln
            m = int(np.ceil(paf_halfwidth)) + 1
            x0, x1 = max(0, int(min(ax, bx)) - m), min(w, int(max(ax, bx)) + m + 1)
            y0, y1 = max(0, int(min(ay, by)) - m), min(h, int(max(ay, by)) + m + 1)
            px, py = xx[y0:y1, x0:x1] - ax, yy[y0:y1, x0:x1] - ay
            along = px * ux + py * uy
            across = np.abs(px * uy - py * ux)
            mask = (along >= -paf_halfwidth) & (along <= ln + paf_halfwidth) & (across <= paf_halfwidth)
            cx, cy = parts + 1 + mapidx[2 * l], parts + 1 + mapidx[2 * l + 1]
            maps[cx, y0:y1, x0:x1][mask] = ux
            maps[cy, y0:y1, x0:x1][mask] = uy
    maps[parts] = 1.0 - maps[:parts].max(axis=0)
    if noise > 0:
        maps += rng.normal(0.0, noise, maps.shape).astype(np.float32)
    return maps, people


def frames(n: int, h: int, w: int, seed: int) -> np.ndarray:
    """n synthetic u8 BGR frames (n, h, w, 3) at network resolution: smooth background + bright blobs."""
    rng = np.random.default_rng(seed)
    out = rng.integers(0, 256, size=(n, h, w, 3), dtype=np.uint8)
    return out


def frames_to_net_input(frames_u8: np.ndarray) -> np.ndarray:
    """u8 HWC BGR -> f32 CHW, x/256 - 0.5 (src/openpose/core/cvMatToOpInput.cpp, uCharCvMatToFloatPtr)."""
    x = frames_u8.astype(np.float32) / np.float32(256.0) - np.float32(0.5)
    return np.ascontiguousarray(x.transpose(0, 3, 1, 2))
