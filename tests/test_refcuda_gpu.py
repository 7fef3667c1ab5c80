"""GPU parity against the reference's OWN CUDA kernels: resizeAndMergeBase.cu, nmsBase.cu, maximumBase.cu,
bodyPartConnectorBase.cu and gpu/cuda.cu compiled unmodified into oracle/_ref/libopref_cuda.so (oracle/Makefile,
`make refcuda`) and run on the same device, same inputs. Everything is compared BIT FOR BIT: the product's CUDA
semantics (OPB200_SEMANTICS_CUDA, the resize and the pre-processing) spell out the FMA contraction nvcc applies to
the reference's sources, so no tolerance is needed -- the tests print the max-abs they saw (0) for the record.
The reference functions only handle one frame correctly (nmsGpu / maximumGpu index N>1 wrongly, the multi-scale
resize reads frame 0), so batches are compared frame by frame."""
import numpy as np
import pytest
import torch

from oracle.post import MODE_CUDA, model_dims, pose_defaults
from openpose_b200 import post
from openpose_b200.op_input import preprocess_gpu
from tests.cases import POST_CASES, assert_peaks_equal, case_id, make_case

pytestmark = pytest.mark.gpu
MAXP = 127


@pytest.fixture(scope="module")
def refcuda(cuda):
    from oracle.ref_cuda import RefCuda, REFCUDA_SO
    import os
    if not os.path.exists(REFCUDA_SO):
        pytest.skip("oracle/_ref/libopref_cuda.so not built (needs /root/reference at build time)")
    return RefCuda()


def _maxabs(a: torch.Tensor, b: torch.Tensor) -> float:
    return float((a - b).abs().max().item())


@pytest.mark.parametrize("shape", [(78, 46, 82), (22, 46, 46), (71, 46, 46), (5, 7, 9), (3, 1, 1), (2, 23, 31)])
def test_resize_x8_equals_reference_kernel(refcuda, cuda, shape):
    g = torch.Generator(device="cpu").manual_seed(1)
    c, h, w = shape
    src = torch.randn((1, c, h, w), generator=g).to(cuda)
    want = torch.empty((1, c, 8 * h, 8 * w), device=cuda)
    refcuda.resize_merge(want, [src])
    got = torch.full_like(want, 7.0)
    post.resize_and_merge_gpu(got, [src])
    torch.cuda.synchronize()
    print("resize x8 %s: max-abs vs resize8TimesKernel = %g" % (shape, _maxabs(got, want)))
    assert torch.equal(got, want)


def test_resize_x8_heatmap_like_values_and_batch(refcuda, cuda):
    """planted-peak maps at net-output resolution (the values the path really sees), two frames."""
    from openpose_b200 import synth
    frames = [synth.planted_maps(0, 46, 82, p, seed=40 + p, sigma=1.2, paf_halfwidth=0.8, noise=0.01,
                                 limb_len=(4.0, 9.0))[0] for p in (3, 20)]
    src = torch.from_numpy(np.stack(frames)).to(cuda)
    got = torch.empty((2, 78, 368, 656), device=cuda)
    post.resize_and_merge_gpu(got, [src])
    for n in range(2):
        want = torch.empty((1, 78, 368, 656), device=cuda)
        refcuda.resize_merge(want, [src[n:n + 1].contiguous()])
        assert torch.equal(got[n], want[0])


@pytest.mark.parametrize("sizes,ratios,target", [
    ([(46, 82), (34, 60), (22, 40), (10, 20)], [1.0, 0.75, 0.5, 0.25], (368, 656)),   # BASELINE config 3
    ([(46, 82), (22, 40)], [1.0, 0.5], (368, 656)),
    ([(8, 12), (6, 9)], [0.5, 0.375], (64, 96)),
    ([(46, 46), (34, 34), (22, 22)], [1.0, 0.75, 0.5], (368, 368)),
])
def test_resize_multiscale_equals_reference_kernel(refcuda, cuda, sizes, ratios, target):
    g = torch.Generator(device="cpu").manual_seed(2)
    c = 78 if target == (368, 656) else 5
    srcs = [torch.randn((1, c, h, w), generator=g).to(cuda) for h, w in sizes]
    want = torch.empty((1, c) + target, device=cuda)
    refcuda.resize_merge(want, srcs, ratios)
    got = torch.full_like(want, 7.0)
    post.resize_and_merge_gpu(got, srcs, ratios)
    torch.cuda.synchronize()
    print("resize %d scales: max-abs vs resizeAndAddAndAverageKernel = %g" % (len(sizes), _maxabs(got, want)))
    assert torch.equal(got, want)


# nmsRegisterKernel's border branch has no x < w guard (nmsBase.cu:86-87): when w is not a multiple of its 16-wide
# blocks, the overhanging threads of row 0 write zeros over the first pixels of row 1, racing with those pixels' own
# threads -- the reference's result is not deterministic there, so only w % 16 == 0 maps are compared.
NMS_CASES = [c for c in POST_CASES if c[2] % 16 == 0]


@pytest.mark.parametrize("case", NMS_CASES, ids=case_id)
def test_nms_cuda_semantics_equals_reference_kernels(refcuda, port, cuda, case):
    parts = model_dims(case[0])[0]
    maps, _ = make_case(case)
    heat = torch.from_numpy(maps[None]).to(cuda)
    want = torch.zeros((1, parts, MAXP + 1, 3), device=cuda)
    refcuda.nms(want, heat, 0.05, (0.25, 0.25))
    got = torch.zeros_like(want)
    post.nms_gpu(got, heat, 0.05, (0.25, 0.25), MODE_CUDA)
    torch.cuda.synchronize()
    assert_peaks_equal(want[0].cpu().numpy(), got[0].cpu().numpy())
    # and the C restatement's mode 1 says the same (keeps the CPU-side oracle honest about the contraction)
    assert_peaks_equal(want[0].cpu().numpy(), port.nms(maps, 0.05, parts, off=(0.25, 0.25), mode=MODE_CUDA))


def test_nms_cuda_semantics_ties_and_small_maps(refcuda, cuda):
    rng = np.random.default_rng(5)
    for h, w in [(8, 16), (5, 32), (4, 16), (12, 48), (3, 16), (33, 128)]:
        m = (rng.integers(0, 4, size=(1, 3, h, w)) / 4.0).astype(np.float32)
        heat = torch.from_numpy(m).to(cuda)
        want = torch.zeros((1, 2, MAXP + 1, 3), device=cuda)
        refcuda.nms(want, heat[:, :2].contiguous(), 0.05, (0.5, 0.5))
        got = torch.zeros_like(want)
        post.nms_gpu(got, heat, 0.05, (0.5, 0.5), MODE_CUDA)
        assert_peaks_equal(want[0].cpu().numpy(), got[0].cpu().numpy())


@pytest.mark.parametrize("case", POST_CASES, ids=case_id)
def test_connect_cuda_semantics_equals_reference(refcuda, cuda, case):
    """pafScoreKernel cube + pafPtrIntoVector + pafVectorIntoPeopleVector + filter + arrays = connectBodyPartsGpu."""
    model = case[0]
    parts = model_dims(model)[0]
    maps, _ = make_case(case)
    heat = torch.from_numpy(maps[None]).to(cuda)
    peaks = torch.zeros((1, parts, MAXP + 1, 3), device=cuda)
    refcuda.nms(peaks, heat, 0.05, (0.25, 0.25))
    for mp in (False, True):
        kp, sc, cube = refcuda.connect(heat[0], peaks[0], model, p=pose_defaults(model, mp), scale=1.959,
                                       maximize_positives=mp)
        params = post.make_connect_params(model, scale_factor=1.959, maximize_positives=mp, semantics=MODE_CUDA)
        got_cube = post.paf_scores_gpu(heat, peaks, params)[0].cpu().numpy()
        assert np.array_equal(got_cube, cube), np.abs(got_cube - cube).max()
        gk, gs, gc = post.connect_body_parts_gpu(heat, peaks, params, capacity=1024)
        n = int(gc[0].item())
        assert n == kp.shape[0], (n, kp.shape)
        assert np.array_equal(gk[0, :n].cpu().numpy(), kp)
        assert np.array_equal(gs[0, :n].cpu().numpy(), sc)


@pytest.mark.parametrize("shape", [(22, 368, 368), (71, 368, 368), (4, 5, 7), (3, 33, 31)])
def test_maximum_equals_reference(refcuda, cuda, shape):
    rng = np.random.default_rng(3)
    c, h, w = shape
    src = rng.normal(0, 1, size=(1, c, h, w)).astype(np.float32)
    src[0, 0] = np.round(src[0, 0])  # ties: thrust::max_element returns the first
    src[0, 1] = 0.25
    s = torch.from_numpy(src).to(cuda)
    want = torch.zeros((1, 1, c - 1, 3), device=cuda)
    refcuda.maximum(want, s)
    got = torch.zeros_like(want)
    post.maximum_gpu(got, s)
    assert torch.equal(got, want)


@pytest.mark.parametrize("hs,ws,net,scale", [
    (120, 160, (368, 656), 3.0666666), (480, 640, (368, 656), 0.76666665), (97, 211, (368, 368), 1.7440758),
    (368, 656, (368, 656), 1.0), (5, 7, (16, 32), 3.2), (720, 1280, (368, 656), 0.51111114),
])
def test_preprocess_equals_reference_kernels(refcuda, cuda, hs, ws, net, scale):
    """reorderAndNormalize + resizeAndPadRbgGpu (the device branch of CvMatToOpInput::createArray)."""
    rng = np.random.default_rng(hs * 1000 + ws)
    frame = rng.integers(0, 256, size=(hs, ws, 3), dtype=np.uint8)
    f = torch.from_numpy(frame).to(cuda)
    want = refcuda.preprocess(f, net, scale)
    got = torch.full((1, 3) + net, 7.0, device=cuda)
    preprocess_gpu(got, f[None].contiguous(), scale)
    torch.cuda.synchronize()
    print("preprocess %dx%d: max-abs vs resizeAndPadKernel = %g" % (hs, ws, _maxabs(got[0], want)))
    assert torch.equal(got[0], want)
