"""Seeded inputs of the golden vectors taken from the reference's own CUDA kernels (tests/golden/make_refcuda_golden.py,
tests/test_golden_refcuda.py). Small on purpose: the fixture is committed."""
import numpy as np

# name, source sizes, scale ratios (None: single scale x8), target, planes, seed
RESIZE_CASES = [
    ("x8_small", [(5, 7)], None, (40, 56), 3, 101),
    ("x8_one_pixel", [(1, 1)], None, (8, 8), 2, 102),
    ("x8_net_like", [(12, 20)], None, (96, 160), 4, 103),
    ("three_scales", [(5, 7), (3, 4), (1, 2)], [1.0, 0.6, 0.2], (40, 56), 5, 104),
    ("four_scales", [(12, 20), (9, 15), (6, 10), (3, 5)], [1.0, 0.75, 0.5, 0.25], (96, 160), 3, 105),
]
# name, parts, h, w, seed   (w % 16 == 0: the reference's nmsRegisterKernel races on other widths, oracle/ref_cuda.py)
NMS_CASES = [("ties_32x48", 3, 32, 48, 201), ("blobs_40x64", 4, 40, 64, 202)]
MAX_CASES = [("small", 5, 17, 23, 301), ("square", 4, 32, 32, 302)]
# name, hs, ws, net (h, w), scale, seed
PRE_CASES = [("up", 12, 20, (48, 64), 3.2, 401), ("down", 45, 80, (24, 40), 0.5, 402), ("pad", 10, 30, (32, 48), 1.5, 403)]


def resize_inputs(sizes, planes, seed):
    rng = np.random.default_rng(seed)
    return [rng.normal(0, 1, size=(1, planes, h, w)).astype(np.float32) for h, w in sizes]


def nms_input(parts, h, w, seed):
    rng = np.random.default_rng(seed)
    if seed % 2:    # quantised values: exact ties everywhere
        return (rng.integers(0, 5, size=(parts + 1, h, w)) / 4.0).astype(np.float32)
    heat = rng.uniform(0, 0.04, size=(parts + 1, h, w)).astype(np.float32)
    yy, xx = np.mgrid[0:h, 0:w]
    for c in range(parts):
        for _ in range(6):
            cy, cx = rng.uniform(0, h), rng.uniform(0, w)
            heat[c] += (0.9 * np.exp(-((yy - cy) ** 2 + (xx - cx) ** 2) / (2 * 2.0 ** 2))).astype(np.float32)
    return heat


def max_input(c, h, w, seed):
    rng = np.random.default_rng(seed)
    heat = rng.normal(0, 1, size=(c, h, w)).astype(np.float32)
    heat[0] = np.round(heat[0])     # ties: the first maximum in raster order wins
    return heat


def pre_input(hs, ws, seed):
    return np.random.default_rng(seed).integers(0, 256, size=(hs, ws, 3), dtype=np.uint8)


# connectBodyPartsGpu: indices into tests.cases.POST_CASES (small maps: the fixture stays small)
CONNECT_CASES = [7, 6]    # BODY_25 46x82 with 2 people, BODY_25 184x328 with 5 people
