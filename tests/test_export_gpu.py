"""GPU parity: opb200_heatmaps_export (through the C ABI) vs the CPU oracle of getHeatMapsCopy -- bit-exact."""
import numpy as np
import pytest
import torch

from openpose_b200 import post
from openpose_b200._lib import Opb200Error
from openpose_b200.pose_extractor import ScaleMode

pytestmark = pytest.mark.gpu


def _heat(n, parts, bkg, pafs, h, w, seed):
    rng = np.random.default_rng(seed)
    m = np.concatenate([rng.uniform(-0.3, 1.3, (n, parts + bkg, h, w)), rng.uniform(-1.4, 1.4, (n, pafs, h, w))],
                       axis=1).astype(np.float32)
    m[:, :, 0, 0] = [0.0, 1.0, -1.0, 0.5][seed % 4]
    m[0, 0, 0, 1] = np.nan                      # fastTruncate passes a NaN through
    m[0, -1, 0, 1] = np.inf
    m[0, 1, 0, 2] = 0.5 / 255.0                 # rounding ties of the UnsignedChar mode
    return m


@pytest.mark.parametrize("mode", list(ScaleMode))
@pytest.mark.parametrize("mask", [1, 2, 4, 3, 5, 6, 7])
def test_export_matches_oracle(port, cuda, mode, mask):
    for (h, w) in [(16, 24), (7, 9)]:          # vector path and scalar path (odd plane size)
        heat = _heat(2, 5, 1, 6, h, w, int(mode) + mask)
        for to_host in (True, False):
            got = post.heatmaps_export_gpu(torch.from_numpy(heat).to(cuda), 5, True, 6, mask, mode, to_host=to_host)
            got = got.cpu().numpy()
            for n in range(2):
                want = port.heatmaps_export(heat[n], 5, True, 6, mask, int(mode))
                assert np.array_equal(want, got[n], equal_nan=True)


def test_export_body25_full_size_into_pinned_memory(port, cuda):
    heat = _heat(1, 25, 1, 52, 368, 656, 3)
    dev = torch.from_numpy(heat).to(cuda)
    out = post.heatmaps_export_gpu(dev, 25, True, 52, 7, ScaleMode.UnsignedChar)
    assert out.is_pinned() and tuple(out.shape) == (1, 78, 368, 656)
    assert np.array_equal(port.heatmaps_export(heat[0], 25, True, 52, 7, int(ScaleMode.UnsignedChar)), out[0].numpy(),
                          equal_nan=True)
    # NoScale with every group selected is a plain copy
    assert np.array_equal(post.heatmaps_export_gpu(dev, 25, True, 52, 7, ScaleMode.NoScale).numpy(), heat,
                          equal_nan=True)
    assert post.heatmaps_export_gpu(dev, 25, True, 52, 0, ScaleMode.NoScale).shape[1] == 0


def test_export_errors(cuda):
    heat = torch.zeros((1, 10, 8, 8), device=cuda)
    with pytest.raises(Opb200Error, match="heatmaps_add_bkg"):
        post.heatmaps_export_gpu(heat, 4, False, 6, 2, ScaleMode.NoScale)
    with pytest.raises(Opb200Error, match="fewer channels"):
        post.heatmaps_export_gpu(heat, 6, True, 6, 7, ScaleMode.NoScale)
    with pytest.raises(Opb200Error, match="ScaleMode"):
        post.heatmaps_export_gpu(heat, 4, True, 4, 7, 9)
