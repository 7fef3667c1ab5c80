"""GPU: the in-library hand / face crop engine (opb200_crop_engine_*: one frame + the affine maps of its crops per batch,
batches dealt round-robin to per-device workers, results in submission order) returns exactly what the separate calls
(opb200_crop_warp -> net -> resize x8 -> maximum) return for the same crops."""
import numpy as np
import pytest
import torch

from openpose_b200 import modelgen, synth
from openpose_b200.crop_extractors import _CropNet, crop_frames_gpu, crop_matrix
from openpose_b200.engine import CropEngine
from openpose_b200._lib import Opb200Error

pytestmark = pytest.mark.gpu


def _rects(rng, n, w, h):
    out = []
    for _ in range(n):
        side = float(rng.uniform(60, 200))
        out.append((float(rng.uniform(-20, w - side + 20)), float(rng.uniform(-20, h - side + 20)), side, side))
    return out


@pytest.mark.parametrize("model,channels", [("hand", 22), ("face", 71)])
def test_crop_engine_matches_step_by_step_calls(cuda, model, channels):
    proto, cm, _, _ = modelgen.build_files(model, 4321)
    w, h, side, batch = 320, 240, 96, 4
    eng = CropEngine(proto, cm, channels, gpus=[0], frame_size=(w, h), net_side=side, batch=batch, lanes=2)
    rng = np.random.default_rng(3)
    frames = [synth.frames(1, h, w, 70 + i)[0] for i in range(4)]
    crops = [4, 1, 3, 4]                                           # crops per frame: full, short batches
    mats = [np.stack([crop_matrix(r, side, mirror=(k % 2 == 1)) for k, r in enumerate(_rects(rng, n, w, h))])
            for n in crops]
    got = []
    eng.submit(100, frames[0], mats[0])
    eng.submit(101, frames[1], mats[1])
    for i in (2, 3):                                               # two in flight (= lanes): take one, submit one
        got.append(eng.next())
        eng.submit(100 + i, frames[i], mats[i])
    while eng.in_flight():
        got.append(eng.next())
    eng.close()
    assert [g[0] for g in got] == [100, 101, 102, 103]
    ref = _CropNet(proto, cm, 0, 8.0)
    ref.init()
    for (fid, peaks), frame, m in zip(got, frames, mats):
        assert peaks.shape == (m.shape[0], channels - 1, 3)
        x = crop_frames_gpu(torch.from_numpy(frame).to(cuda), list(m), (side, side))
        want = ref.run(x)
        assert np.array_equal(peaks, want)
    ref.net.close()


def test_crop_engine_rejects_bad_use(cuda):
    proto, cm, _, _ = modelgen.build_files("hand", 4321)
    with pytest.raises(Opb200Error, match="multiple of 8"):
        CropEngine(proto, cm, 22, frame_size=(320, 240), net_side=100, batch=2)
    eng = CropEngine(proto, cm, 22, frame_size=(320, 240), net_side=96, batch=2)
    frame = synth.frames(1, 240, 320, 1)[0]
    with pytest.raises(Opb200Error, match="between 1 and `batch` crops"):
        eng.submit(0, frame, np.zeros((3, 2, 3)))
    with pytest.raises(Opb200Error, match="nothing was submitted"):
        eng.next()
    eng.close()
