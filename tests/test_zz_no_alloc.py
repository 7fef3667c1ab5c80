"""opb200_nms_static (the entry point that allocates nothing: scratch in a static device array inside the library)
against opb200_nms, through the C ABI, on two calls in a row. Collected last on purpose (file name).
Passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest
import torch

from openpose_b200 import _lib, post
from tests.cases import POST_CASES, make_case

pytestmark = pytest.mark.gpu


def _nms_static(peaks, heat, thr, off, mode):
    L = _lib.lib()
    rc = L.opb200_nms_static(peaks.data_ptr(), heat.data_ptr(), float(thr), float(off[0]), float(off[1]),
                             _lib.i4(peaks.shape), _lib.i4(heat.shape), int(mode),
                             torch.cuda.current_stream().cuda_stream)
    _lib.check(rc)
    return peaks


def _equal_up_to_counts(a, b):
    a, b = a.cpu().numpy(), b.cpu().numpy()
    cnt = a[..., 0, 0].astype(int)
    assert np.array_equal(cnt, b[..., 0, 0].astype(int))
    fa, fb = a.reshape(-1, a.shape[-2], 3), b.reshape(-1, b.shape[-2], 3)
    for c, n in enumerate(cnt.reshape(-1)):
        assert np.array_equal(fa[c, : n + 1], fb[c, : n + 1])   # bit-exact: the tolerance of the nms tests


def test_nms_static_matches_nms_twice_in_a_row(cuda):
    # call 1: a crowd frame (many peaks); call 2: two sparse frames at another size -- the shared array keeps the
    # first call's contents, so the second call checks that stale scratch is harmless
    dense = torch.from_numpy(make_case(POST_CASES[3])[0][None]).to(cuda)                       # (1,78,368,656)
    sparse = torch.from_numpy(np.stack([make_case(POST_CASES[6])[0]] * 2)).to(cuda)            # (2,78,184,328)
    sparse[1] = sparse[1].flip(-1)
    for heat in (dense, sparse):
        for mode in (_lib.SEMANTICS_CPU, _lib.SEMANTICS_CUDA):
            want = post.nms_gpu(torch.zeros((heat.shape[0], 25, 128, 3), device=cuda), heat, 0.05, (0.25, 0.25), mode)
            got = _nms_static(torch.zeros_like(want), heat, 0.05, (0.25, 0.25), mode)
            torch.cuda.synchronize()
            _equal_up_to_counts(want, got)
            assert int(got[:, :, 0, 0].sum().item()) > 0


def test_nms_static_rejects_shapes_beyond_the_static_array(cuda):
    heat = torch.zeros((8, 78, 46, 82), device=cuda)         # 8 BODY_25 frames need 6.5 MB > 4 MiB
    peaks = torch.zeros((8, 25, 128, 3), device=cuda)
    assert _lib.lib().opb200_nms_scratch_bytes(_lib.i4(peaks.shape), _lib.i4(heat.shape)) > (4 << 20)
    with pytest.raises(_lib.Opb200Error, match="OPB200_NMS_STATIC_SCRATCH_BYTES"):
        _nms_static(peaks, heat, 0.05, (0.5, 0.5), 0)
