"""GPU parity: CUDA post-processing (through the C ABI) vs the CPU oracle -- bit-exact."""
import numpy as np
import pytest
import torch

from oracle.post import MODE_CPU, MODE_CUDA, model_dims
from openpose_b200 import post, synth
from openpose_b200._lib import Opb200Error
from tests.cases import POST_CASES, assert_peaks_equal, case_id, make_case

pytestmark = pytest.mark.gpu
MAXP = 127


def _nms_gpu(maps, parts, thr, off, mode, cuda, stale=None):
    heat = torch.from_numpy(maps).to(cuda)
    n = heat.shape[0]
    peaks = torch.zeros((n, parts, MAXP + 1, 3), device=cuda) if stale is None else stale
    post.nms_gpu(peaks, heat, thr, off, mode)
    return peaks


@pytest.mark.parametrize("case", POST_CASES, ids=case_id)
@pytest.mark.parametrize("mode", [MODE_CPU, MODE_CUDA])
def test_nms_matches_oracle(port, cuda, case, mode):
    parts = model_dims(case[0])[0]
    maps, _ = make_case(case)
    want = port.nms(maps, 0.05, parts, off=(0.25, 0.25), mode=mode)
    got = _nms_gpu(maps[None], parts, 0.05, (0.25, 0.25), mode, cuda)[0].cpu().numpy()
    assert_peaks_equal(want, got)


def test_nms_ties_small_maps_and_batch(port, cuda):
    rng = np.random.default_rng(5)
    for h, w in [(8, 9), (5, 5), (4, 7), (12, 6), (3, 3), (2, 5), (33, 129)]:
        m = (rng.integers(0, 4, size=(6, 3, h, w)) / 4.0).astype(np.float32)  # batch of 6 frames, exact ties
        for mode in (MODE_CPU, MODE_CUDA):
            got = _nms_gpu(m, 2, 0.05, (0.5, 0.5), mode, cuda).cpu().numpy()
            for n in range(6):
                assert_peaks_equal(port.nms(m[n], 0.05, 2, mode=mode), got[n])


def test_nms_leaves_stale_rows_and_validates_threshold(port, cuda):
    maps, _ = make_case(POST_CASES[0])
    stale = torch.full((1, 25, MAXP + 1, 3), 7.0, device=cuda)
    got = _nms_gpu(maps[None], 25, 0.05, (0.5, 0.5), MODE_CPU, cuda, stale=stale)[0].cpu().numpy()
    for c in range(25):
        n = int(got[c, 0, 0])
        assert (got[c, n + 1:] == 7.0).all()
    with pytest.raises(Opb200Error, match="threshold value invalid"):
        _nms_gpu(maps[None], 25, 1.5, (0.5, 0.5), MODE_CPU, cuda)


@pytest.mark.parametrize("case", POST_CASES, ids=case_id)
def test_connect_matches_oracle(port, cuda, case):
    model = case[0]
    parts = model_dims(model)[0]
    maps, _ = make_case(case)
    peaks = port.nms(maps, 0.05, parts, off=(0.25, 0.25))
    heat_d = torch.from_numpy(maps[None]).to(cuda)
    peaks_d = torch.from_numpy(peaks[None]).to(cuda)
    for mp in (False, True):
        kp, sc, rows = port.connect(maps, peaks, model, scale=1.959, maximize_positives=mp)
        params = post.make_connect_params(model, scale_factor=1.959, maximize_positives=mp)
        gk, gs, gc = post.connect_body_parts_gpu(heat_d, peaks_d, params, capacity=1024)
        n = int(gc[0].item())
        assert n == kp.shape[0], (n, kp.shape)
        assert np.array_equal(gk[0, :n].cpu().numpy(), kp)
        assert np.array_equal(gs[0, :n].cpu().numpy(), sc)
    for mode in (MODE_CPU, MODE_CUDA):
        params = post.make_connect_params(model, semantics=mode)
        cube = post.paf_scores_gpu(heat_d, peaks_d, params)[0].cpu().numpy()
        assert np.array_equal(cube, port.pair_scores(maps, peaks, model, mode=mode))


@pytest.mark.parametrize("case", POST_CASES, ids=case_id)
def test_connect_cuda_semantics_matches_reference(port, cuda, case):
    """CUDA-build path: CUDA-rule score cube -> pafPtrIntoVector (one global sort) -> pafVectorIntoPeopleVector
    (five-case assignment incl. merging) -> filter -> arrays; against the reference's own object code when
    oracle/_ref is built, else the C restatement (tests/test_oracle_vs_ref.py keeps the two bit-equal)."""
    from oracle.post import Ref
    impl = Ref() if Ref.available() else port
    model = case[0]
    parts = model_dims(model)[0]
    maps, _ = make_case(case)
    peaks = port.nms(maps, 0.05, parts, off=(0.25, 0.25))
    heat_d = torch.from_numpy(maps[None]).to(cuda)
    peaks_d = torch.from_numpy(peaks[None]).to(cuda)
    from oracle.post import pose_defaults
    for mp in (False, True):
        cube = port.pair_scores(maps, peaks, model, p=pose_defaults(model, mp), mode=MODE_CUDA)
        want = impl.connect_from_pair_scores(cube, peaks, model, scale=1.959, maximize_positives=mp)
        kp, sc = want[0], want[1]
        params = post.make_connect_params(model, scale_factor=1.959, maximize_positives=mp, semantics=MODE_CUDA)
        gk, gs, gc = post.connect_body_parts_gpu(heat_d, peaks_d, params, capacity=1024)
        n = int(gc[0].item())
        assert n == kp.shape[0], (n, kp.shape)
        assert np.array_equal(gk[0, :n].cpu().numpy(), kp)
        assert np.array_equal(gs[0, :n].cpu().numpy(), sc)


def test_connect_maximum_number_of_candidates(port, cuda):
    """A uniform PAF field pointing to +x with even parts on the left half and odd parts on the right: every pair of a
    limb that runs even -> odd passes the line integral: 127 x 127 = 16129 candidates, the most a limb can have (full
    candidate lists of pafCandKernel, full 16384-key sort). Result must equal the oracle's std::sort + greedy matching."""
    h, w, parts, n = 96, 160, 25, 127
    rng = np.random.default_rng(17)
    maps = np.zeros((78, h, w), np.float32)
    maps[26::2] = 0.5                                   # x components of every limb's PAF
    peaks = np.zeros((parts, 128, 3), np.float32)
    for p in range(parts):
        lo, hi = (2, w // 2 - 8) if p % 2 == 0 else (w // 2 + 8, w - 3)
        peaks[p, 0, 0] = n
        peaks[p, 1:n + 1, 0] = rng.uniform(lo, hi, n)
        peaks[p, 1:n + 1, 1] = rng.uniform(2, h - 3, n)
        peaks[p, 1:n + 1, 2] = rng.uniform(0.1, 1.0, n)
    cube = port.pair_scores(maps, peaks, 0, mode=MODE_CPU)
    assert int((cube > 1e-6).sum(axis=(1, 2)).max()) > 8192
    kp, sc, _ = port.connect(maps, peaks, 0, scale=1.0)
    params = post.make_connect_params(0, scale_factor=1.0)
    gk, gs, gc = post.connect_body_parts_gpu(torch.from_numpy(maps[None]).to(cuda), torch.from_numpy(peaks[None]).to(cuda),
                                             params, capacity=4096)
    n = int(gc[0].item())
    assert n == kp.shape[0] and n > 0
    assert np.array_equal(gk[0, :n].cpu().numpy(), kp) and np.array_equal(gs[0, :n].cpu().numpy(), sc)


def test_connect_cuda_semantics_batch(port, cuda):
    cases = [POST_CASES[1], POST_CASES[5], POST_CASES[3]]
    maps = np.stack([make_case(c)[0] for c in cases])
    peaks = np.stack([port.nms(m, 0.05, 25, off=(0.25, 0.25)) for m in maps])
    params = post.make_connect_params(0, scale_factor=2.0, semantics=MODE_CUDA)
    gk, gs, gc = post.connect_body_parts_gpu(torch.from_numpy(maps).to(cuda), torch.from_numpy(peaks).to(cuda),
                                             params, capacity=256)
    for i in range(len(cases)):
        cube = port.pair_scores(maps[i], peaks[i], 0, mode=MODE_CUDA)
        kp, sc = port.connect_from_pair_scores(cube, peaks[i], 0, scale=2.0)[:2]
        n = int(gc[i].item())
        assert n == kp.shape[0]
        assert np.array_equal(gk[i, :n].cpu().numpy(), kp) and np.array_equal(gs[i, :n].cpu().numpy(), sc)


def test_connect_batch_of_frames(port, cuda):
    cases = [POST_CASES[1], POST_CASES[2], POST_CASES[5], POST_CASES[0]]
    maps = np.stack([make_case(c)[0] for c in cases])
    peaks = np.stack([port.nms(m, 0.05, 25, off=(0.25, 0.25)) for m in maps])
    params = post.make_connect_params(0, scale_factor=2.0)
    gk, gs, gc = post.connect_body_parts_gpu(torch.from_numpy(maps).to(cuda), torch.from_numpy(peaks).to(cuda),
                                             params, capacity=256)
    for i in range(len(cases)):
        kp, sc, _ = port.connect(maps[i], peaks[i], 0, scale=2.0)
        n = int(gc[i].item())
        assert n == kp.shape[0]
        assert np.array_equal(gk[i, :n].cpu().numpy(), kp) and np.array_equal(gs[i, :n].cpu().numpy(), sc)


@pytest.mark.parametrize("shape", [(78, 46, 82), (22, 46, 46), (5, 7, 9), (3, 1, 1), (2, 23, 31)])
def test_resize_x8_matches_oracle(port, cuda, shape):
    rng = np.random.default_rng(1)
    c, h, w = shape
    src = rng.normal(0, 1, size=(2, c, h, w)).astype(np.float32)
    dst = torch.empty((2, c, 8 * h, 8 * w), device=cuda)
    post.resize_and_merge_gpu(dst, [torch.from_numpy(src).to(cuda)])
    got = dst.cpu().numpy()
    for n in range(2):
        assert np.array_equal(got[n], port.resize_merge([src[n]], 8 * h, 8 * w))


def test_resize_multiscale_matches_oracle(port, cuda):
    """config 3: scale_number 4, scale_gap 0.25 -> net outputs 82x46, 60x34, 40x22, 20x10."""
    rng = np.random.default_rng(2)
    sizes = [(46, 82), (34, 60), (22, 40), (10, 20)]
    ratios = [1.0, 0.75, 0.5, 0.25]
    srcs = [rng.normal(0, 1, size=(1, 78, h, w)).astype(np.float32) for h, w in sizes]
    dst = torch.empty((1, 78, 368, 656), device=cuda)
    post.resize_and_merge_gpu(dst, [torch.from_numpy(s).to(cuda) for s in srcs], ratios)
    want = port.resize_merge([s[0] for s in srcs], 368, 656, ratios)
    assert np.array_equal(dst[0].cpu().numpy(), want)


@pytest.mark.parametrize("sizes,ratios,target,planes", [
    ([(5, 7), (3, 4), (1, 2)], [1.0, 0.6, 0.2], (40, 56), 5),          # tiny / one-row sources, three scales
    ([(2, 3), (2, 2)], [1.0, 0.75], (16, 24), 3),                      # every tap clamps
    ([(40, 30), (29, 22)], [1.0, 0.73], (330, 245), 2),                # odd target width, several row chunks
    ([(8, 8), (64, 64)], [1.0, 16.0], (16, 16), 3),                    # second scale down-samples: band kernel
    ([(6, 6)] * 5, [1.0, 0.9, 0.8, 0.7, 0.6], (48, 48), 2),            # five scales: band kernel
])
def test_resize_multiscale_ragged_matches_oracle(port, cuda, sizes, ratios, target, planes):
    """the streaming multi-scale kernel (2-4 up-sampling scales) and the band kernel it falls back to, on ragged shapes."""
    rng = np.random.default_rng(11)
    srcs = [rng.normal(0, 1, size=(1, planes, h, w)).astype(np.float32) for h, w in sizes]
    dst = torch.full((1, planes) + target, 7.0, device=cuda)
    post.resize_and_merge_gpu(dst, [torch.from_numpy(s).to(cuda) for s in srcs], ratios)
    want = port.resize_merge([s[0] for s in srcs], target[0], target[1], ratios)
    assert np.array_equal(dst[0].cpu().numpy(), want)


@pytest.mark.parametrize("guard", [64, 1])
@pytest.mark.parametrize("sizes,ratios,target,planes", [
    ([(46, 82)], None, (368, 656), 3),                                  # x8 streaming kernel (aligned) / band kernel
    ([(5, 7)], None, (40, 56), 5),
    ([(46, 82), (34, 60), (22, 40), (10, 20)], [1.0, 0.75, 0.5, 0.25], (368, 656), 2),
    ([(5, 7), (3, 4), (1, 2)], [1.0, 0.6, 0.2], (40, 56), 5),
])
def test_resize_writes_only_its_target(port, cuda, sizes, ratios, target, planes, guard):
    """out-of-bounds canary (the pool has no compute-sanitizer): the target is a window of a larger buffer whose
    margins hold a sentinel; guard = 1 also takes the target off 16-byte alignment (the non-vectorised kernels)."""
    rng = np.random.default_rng(5)
    srcs = [rng.normal(0, 1, size=(1, planes, h, w)).astype(np.float32) for h, w in sizes]
    n = planes * target[0] * target[1]
    big = torch.full((n + 2 * guard,), -12345.0, device=cuda)
    dst = big[guard:guard + n].view(1, planes, target[0], target[1])
    post.resize_and_merge_gpu(dst, [torch.from_numpy(s).to(cuda) for s in srcs], ratios)
    torch.cuda.synchronize()
    assert bool((big[:guard] == -12345.0).all()) and bool((big[guard + n:] == -12345.0).all())
    want = port.resize_merge([s[0] for s in srcs], target[0], target[1], ratios)
    assert np.array_equal(dst[0].cpu().numpy(), want)


def test_resize_errors_and_copy(cuda):
    src = torch.randn((1, 3, 10, 12), device=cuda)
    dst = torch.empty((1, 3, 10, 12), device=cuda)
    post.resize_and_merge_gpu(dst, [src])
    assert torch.equal(dst, src)
    with pytest.raises(Opb200Error, match="8x resize"):
        post.resize_and_merge_gpu(torch.empty((1, 3, 40, 48), device=cuda), [src])
    with pytest.raises(Opb200Error, match="More than 8 scales"):
        post.resize_and_merge_gpu(torch.empty((1, 3, 80, 96), device=cuda), [src] * 9, [1.0] * 9)


@pytest.mark.parametrize("shape", [(22, 368, 368), (71, 368, 368), (4, 5, 7), (3, 33, 31)])
def test_maximum_matches_oracle(port, cuda, shape):
    rng = np.random.default_rng(3)
    c, h, w = shape
    src = rng.normal(0, 1, size=(3, c, h, w)).astype(np.float32)
    src[0, 0] = np.round(src[0, 0])  # ties: first occurrence in raster order must win
    src[1, 1] = 0.25
    out = torch.empty((3, 1, c - 1, 3), device=cuda)
    post.maximum_gpu(out, torch.from_numpy(src).to(cuda))
    got = out.cpu().numpy()
    for n in range(3):
        assert np.array_equal(got[n, 0], port.maximum(src[n], c - 1))


def test_full_size_properties(cuda):
    """BASELINE-size property checks that need no oracle: idempotence of nms on its own peak map,
    resize of a constant is that constant, maximum of a planted spike."""
    const = torch.full((1, 78, 46, 82), 0.375, device=cuda)
    dst = torch.empty((1, 78, 368, 656), device=cuda)
    post.resize_and_merge_gpu(dst, [const])
    assert torch.equal(dst, torch.full_like(dst, 0.375))
    hm = torch.zeros((4, 71, 368, 368), device=cuda)
    ys = torch.randint(0, 368, (4, 70))
    xs = torch.randint(0, 368, (4, 70))
    for n in range(4):
        hm[n, torch.arange(70), ys[n], xs[n]] = 0.9
    out = torch.empty((4, 1, 70, 3), device=cuda)
    post.maximum_gpu(out, hm)
    assert torch.equal(out[:, 0, :, 0].cpu(), xs.float()) and torch.equal(out[:, 0, :, 1].cpu(), ys.float())
