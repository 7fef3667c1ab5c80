"""GPU, needs two devices (skipped on a one-GPU box): the in-library engines deal batches round the devices and hand the
results back in submission order -- same bits as one device, whatever device ran a batch."""
import numpy as np
import pytest
import torch

from openpose_b200 import modelgen, synth
from openpose_b200.crop_extractors import crop_matrix
from openpose_b200.engine import CropEngine, PoseEngine

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs")]
LOOSE = dict(nms_threshold=0.02, inter_threshold=0.0, inter_min_above=0.3, min_subset_cnt=2, min_subset_score=0.0)


def _run_pose(gpus, proto, cm, batches):
    eng = PoseEngine(proto, cm, gpus=gpus, net_size=(160, 96), frame_size=(320, 180), batch=3, lanes=2, capacity=64,
                     overrides=LOOSE)
    got, fid, cap = [], 0, 2 * len(gpus)                      # lanes x devices batches may be in flight
    for b in batches:
        if eng.in_flight() == cap:
            got.append(eng.next())
        eng.submit(fid, b)
        fid += b.shape[0]
    while eng.in_flight():
        got.append(eng.next())
    eng.close()
    return got


def test_pose_engine_two_devices_equals_one(cuda):
    proto, cm, _, _ = modelgen.build_files("body_25", 1234)
    batches = [synth.frames(3 if i != 4 else 2, 180, 320, 90 + i) for i in range(9)]
    one = _run_pose([0], proto, cm, batches)
    two = _run_pose([0, 1], proto, cm, batches)
    assert [g[0] for g in two] == [g[0] for g in one] == [0, 3, 6, 9, 12, 14, 17, 20, 23]
    people = 0
    for (_, kp1, sc1, c1), (_, kp2, sc2, c2) in zip(one, two):
        assert np.array_equal(c1, c2)
        for f in range(c1.shape[0]):
            n = min(int(c1[f]), 64)
            people += n
            assert np.array_equal(kp1[f, :n], kp2[f, :n]) and np.array_equal(sc1[f, :n], sc2[f, :n])
    assert people > 0


def test_crop_engine_two_devices_equals_one(cuda):
    proto, cm, _, _ = modelgen.build_files("hand", 4321)
    rng = np.random.default_rng(8)
    frames = [synth.frames(1, 240, 320, 60 + i)[0] for i in range(7)]
    mats = []
    for i in range(7):
        n = int(rng.integers(1, 5))
        mats.append(np.stack([crop_matrix((float(rng.uniform(0, 200)), float(rng.uniform(0, 120)), 100.0, 100.0), 96,
                                          mirror=bool(k % 2)) for k in range(n)]))
    res = []
    for gpus in ([0], [0, 1]):
        eng = CropEngine(proto, cm, 22, gpus=gpus, frame_size=(320, 240), net_side=96, batch=4, lanes=2)
        got = []
        for i in range(7):
            if eng.in_flight() == 2 * len(gpus):
                got.append(eng.next())
            eng.submit(i, frames[i], mats[i])
        while eng.in_flight():
            got.append(eng.next())
        eng.close()
        res.append(got)
    assert [g[0] for g in res[0]] == [g[0] for g in res[1]] == list(range(7))
    for (_, a), (_, b), m in zip(res[0], res[1], mats):
        assert a.shape == (m.shape[0], 21, 3) and np.array_equal(a, b)
