"""Generates tests/golden/refcuda_golden.npz from the REFERENCE's own CUDA kernels (oracle/_ref/libopref_cuda.so, built from
/root/reference's .cu files by oracle/Makefile and shipped to the GPU box): resizeAndMergeGpu (x8 and three scales), nmsGpu,
maximumGpu, reorderAndNormalize + resizeAndPadRbgGpu, connectBodyPartsGpu's pair-score cube. Needs a GPU:
    gpurun -- 'python tests/golden/make_refcuda_golden.py gpurun_out/refcuda_golden.npz'
Inputs are re-generated from the seeds in tests/cases_refcuda.py; a CRC of every input is stored to catch generator drift."""
import os
import sys
import zlib

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.ref_cuda import RefCuda  # noqa: E402
from tests import cases_refcuda as K  # noqa: E402


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "tests", "golden", "refcuda_golden.npz")
    ref, dev, out = RefCuda(), torch.device("cuda:0"), {}
    crc = lambda a: np.array([zlib.crc32(np.ascontiguousarray(a).tobytes())], np.uint32)  # noqa: E731
    for name, sizes, ratios, target, planes, seed in K.RESIZE_CASES:
        srcs = K.resize_inputs(sizes, planes, seed)
        dst = torch.empty((1, planes) + target, device=dev)
        ref.resize_merge(dst, [torch.from_numpy(s).to(dev) for s in srcs], ratios)
        out["resize/%s/crc" % name] = crc(np.concatenate([s.ravel() for s in srcs]))
        out["resize/%s/out" % name] = dst[0].cpu().numpy()
    for name, parts, h, w, seed in K.NMS_CASES:
        heat = K.nms_input(parts, h, w, seed)
        peaks = torch.zeros((1, parts, 128, 3), device=dev)
        ref.nms(peaks, torch.from_numpy(heat[None]).to(dev), 0.05, (0.5, 0.5))
        p = peaks[0].cpu().numpy()
        for c in range(parts):                         # zero the stale rows so the fixture is deterministic
            p[c, int(p[c, 0, 0]) + 1:] = 0
            p[c, 0, 1:] = 0
        out["nms/%s/crc" % name] = crc(heat)
        out["nms/%s/peaks" % name] = p
    for name, c, h, w, seed in K.MAX_CASES:
        heat = K.max_input(c, h, w, seed)
        tgt = torch.zeros((1, 1, c - 1, 3), device=dev)
        ref.maximum(tgt, torch.from_numpy(heat[None]).to(dev))
        out["max/%s/crc" % name] = crc(heat)
        out["max/%s/out" % name] = tgt[0, 0].cpu().numpy()
    for name, hs, ws, net, scale, seed in K.PRE_CASES:
        frame = K.pre_input(hs, ws, seed)
        res = ref.preprocess(torch.from_numpy(frame).to(dev), net, scale)
        out["pre/%s/crc" % name] = crc(frame)
        out["pre/%s/out" % name] = res.cpu().numpy()
    from oracle.post import MODE_CUDA, Port, model_dims, pose_defaults
    from tests.cases import POST_CASES, case_id, make_case
    port = Port()
    for ci in K.CONNECT_CASES:
        case = POST_CASES[ci]
        model = case[0]
        parts = model_dims(model)[0]
        maps, _ = make_case(case)
        peaks = port.nms(maps, 0.05, parts, off=(0.25, 0.25), mode=MODE_CUDA)      # the same peaks on both sides
        cid = case_id(case)
        out["connect/%s/crc" % cid] = crc(maps)
        for mp in (0, 1):
            kp, sc, cube = ref.connect(torch.from_numpy(maps).to(dev), torch.from_numpy(peaks).to(dev), model,
                                       p=pose_defaults(model, bool(mp)), scale=1.959, maximize_positives=bool(mp))
            out["connect/%s/kp%d" % (cid, mp)] = kp
            out["connect/%s/sc%d" % (cid, mp)] = sc
            out["connect/%s/cube%d" % (cid, mp)] = cube        # mostly zeros: compresses to a few KB
    path = out_path
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB,", len(out), "arrays")


if __name__ == "__main__":
    main()
