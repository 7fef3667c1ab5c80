"""GPU parity: opb200_preprocess (through the C ABI) vs the CPU oracle -- bit-exact."""
import numpy as np
import pytest
import torch

from openpose_b200._lib import Opb200Error
from openpose_b200.op_input import CvMatToOpInputB200, ScaleAndSizeExtractor, preprocess_gpu

pytestmark = pytest.mark.gpu


def _frames(n, h, w, seed):
    rng = np.random.default_rng(seed)
    f = rng.integers(0, 256, size=(n, h, w, 3), dtype=np.uint8)
    f[:, : h // 3] = (np.arange(w)[None, None, :, None] * 7 % 256).astype(np.uint8)  # a smooth band as well as noise
    return f


@pytest.mark.parametrize("hs,ws,net", [
    (120, 160, (656, 368)),     # up-sampling, width-limited padding on the bottom
    (480, 640, (656, 368)),     # down-sampling, padding on the right
    (97, 211, (368, 368)),      # odd sizes
    (368, 656, (656, 368)),     # scale 1
    (5, 7, (32, 16)),           # tiny source: every tap clamps
    (300, 40, (16, 304)),       # tall
])
def test_preprocess_matches_oracle(port, cuda, hs, ws, net):
    frames = _frames(3, hs, ws, hs * 1000 + ws)
    scales, sizes, _, _ = ScaleAndSizeExtractor(net).extract((ws, hs))
    got = CvMatToOpInputB200(device=cuda).createArray(frames, scales, sizes)[0].cpu().numpy()
    assert got.shape == (3, 3, net[1], net[0])
    for n in range(3):
        want = port.preprocess(frames[n], (net[1], net[0]), scales[0])
        assert np.array_equal(want, got[n]), (n, np.abs(want - got[n]).max())


def test_preprocess_multi_scale_and_arbitrary_scale(port, cuda):
    frames = _frames(1, 240, 320, 9)
    scales, sizes, _, _ = ScaleAndSizeExtractor((656, 368), -1.0, (-1, -1), 3, 0.25).extract((320, 240))
    outs = CvMatToOpInputB200(device=cuda).createArray(frames[0], scales, sizes)
    assert [tuple(o.shape) for o in outs] == [(1, 3, h, w) for (w, h) in sizes]
    for o, s, (w, h) in zip(outs, scales, sizes):
        assert np.array_equal(port.preprocess(frames[0], (h, w), s), o[0].cpu().numpy())
    # scales that put the padding edge at awkward places
    src = torch.from_numpy(frames).to(cuda)
    for s in (0.3333333, 1.0, 1.7, 2.049):
        dst = torch.full((1, 3, 368, 656), 7.0, device=cuda)
        preprocess_gpu(dst, src, s)
        assert np.array_equal(port.preprocess(frames[0], (368, 656), s), dst[0].cpu().numpy())


def test_preprocess_full_size_properties(cuda):
    """1080p frames at the bench size: scale-1 identity, padding zero, value range."""
    frames = torch.from_numpy(_frames(4, 1080, 1920, 1)).to(cuda)
    dst = torch.empty((4, 3, 1088, 1920), device=cuda)
    preprocess_gpu(dst, frames, 1.0)
    want = frames.permute(0, 3, 1, 2).float() / 256.0 - 0.5
    assert torch.equal(dst[:, :, :1080], want)
    assert not dst[:, :, 1080:].any()
    scales, sizes, _, _ = ScaleAndSizeExtractor((-1, 368)).extract((1920, 1080))
    assert sizes == [(656, 368)]
    out = CvMatToOpInputB200(device=cuda).createArray(frames, scales, sizes)[0]
    edge = int(np.ceil(np.float32(1080) * np.float32(scales[0])))
    assert out.shape == (4, 3, 368, 656)
    # Catmull-Rom weights have L1 norm <= 1.25 per axis, so away from the extrapolating last rows / columns
    # (the reference does not clamp the fractional offset there) |value| <= 0.5 * 1.25^2
    assert out[:, :, : edge - 3, :648].abs().max() <= 0.5 * 1.5625 + 1e-5
    assert not out[:, :, edge:].any() and out[:, :, edge - 1].any()


def test_preprocess_errors(cuda):
    x = CvMatToOpInputB200(device=cuda)
    with pytest.raises(Opb200Error, match="empty inputData"):
        x.createArray(np.zeros((0, 10, 3), np.uint8), [1.0], [(16, 16)])
    with pytest.raises(Opb200Error, match="3-channel BGR"):
        x.createArray(np.zeros((10, 10, 4), np.uint8), [1.0], [(16, 16)])
    with pytest.raises(Opb200Error, match="size"):
        x.createArray(np.zeros((10, 10, 3), np.uint8), [1.0, 0.5], [(16, 16)])
    with pytest.raises(Opb200Error):
        CvMatToOpInputB200(gpuResize=False)
    with pytest.raises(Opb200Error):
        preprocess_gpu(torch.empty((1, 3, 16, 16), device=cuda), torch.zeros((1, 4, 4, 3), dtype=torch.uint8,
                                                                             device=cuda), 0.0)
