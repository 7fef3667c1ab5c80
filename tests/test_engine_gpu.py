"""GPU: the in-library engine (opb200_engine_*: worker thread + compute / copy streams + pinned staging per device, batches
dealt round-robin, results in frame order) returns exactly what the same C-ABI calls return when made one by one."""
import numpy as np
import pytest
import torch

from openpose_b200 import modelgen, post, synth
from openpose_b200.engine import PoseEngine
from openpose_b200.net import NetB200
from openpose_b200.op_input import preprocess_gpu

pytestmark = pytest.mark.gpu
# thresholds loose enough that the noise maps of a random-init net assemble "people": every stage then carries data
LOOSE = dict(nms_threshold=0.02, inter_threshold=0.0, inter_min_above=0.3, min_subset_cnt=2, min_subset_score=0.0)


def _reference_chain(proto, cm, frames, eng, cuda):
    """the engine's chain through the separate entry points"""
    net = NetB200(proto, cm)
    net.initializationOnThread()
    n = frames.shape[0]
    x = torch.empty((n, 3, eng.cfg.net_h, eng.cfg.net_w), device=cuda)
    preprocess_gpu(x, torch.from_numpy(frames).to(cuda), eng.cfg.scale_input_to_net)
    net.forwardPass(x)
    out = net.getOutputBlobArray()
    heat = torch.empty((n, out.shape[1], eng.cfg.net_h, eng.cfg.net_w), device=cuda)
    post.resize_and_merge_gpu(heat, [out])
    peaks = torch.zeros((n, eng.parts, eng.cfg.connect.max_peaks + 1, 3), device=cuda)
    off = float(np.float32(0.5 / float(eng.cfg.connect.scale_factor)))
    post.nms_gpu(peaks, heat, eng.cfg.nms_threshold, (off, off))
    kp, sc, cnt = post.connect_body_parts_gpu(heat, peaks, eng.cfg.connect, capacity=eng.capacity)
    res = kp.cpu().numpy(), sc.cpu().numpy(), cnt.cpu().numpy()
    net.close()
    return res


@pytest.mark.parametrize("model,pose_model", [("body_25", 0), ("coco", 1)])
def test_engine_matches_step_by_step_calls_and_keeps_frame_order(cuda, model, pose_model):
    proto, cm, _, _ = modelgen.build_files(model, 1234)
    frame_size, net_size, batch = (320, 180), (160, 96), 3
    eng = PoseEngine(proto, cm, gpus=[0], pose_model=pose_model, net_size=net_size, frame_size=frame_size, batch=batch,
                     lanes=2, capacity=64, overrides=LOOSE)
    batches = [synth.frames(batch, frame_size[1], frame_size[0], 50 + i) for i in range(5)]
    batches[3] = batches[3][:2]                                   # a short batch in the middle
    fid = 0
    for b in batches[:2]:
        eng.submit(fid, b)
        fid += b.shape[0]
    got = []
    for i, b in enumerate(batches[2:]):                           # keep two in flight (= lanes): take one, submit one
        got.append(eng.next())
        staging = eng.acquire()                                   # (with every lane busy this would wait for a next())
        staging[:b.shape[0]] = b
        eng.commit(fid, b.shape[0])
        fid += b.shape[0]
    while eng.in_flight():
        got.append(eng.next())
    assert [g[0] for g in got] == [0, 3, 6, 9, 11]                # submission order = frame order
    people = 0
    for (first, kp, sc, cnt), b in zip(got, batches):
        assert kp.shape[0] == b.shape[0]
        wk, ws, wc = _reference_chain(proto, cm, b, eng, cuda)
        assert np.array_equal(cnt, wc)
        for f in range(b.shape[0]):
            n = min(int(cnt[f]), eng.capacity)
            people += n
            assert np.array_equal(kp[f, :n], wk[f, :n]) and np.array_equal(sc[f, :n], ws[f, :n])
    assert people > 0
    eng.close()


def test_engine_rejects_bad_configuration(cuda):
    from openpose_b200._lib import Opb200Error
    proto, cm, _, _ = modelgen.build_files("body_25", 1234)
    with pytest.raises(Opb200Error, match="multiples of 16"):          # ScaleAndSizeExtractor's own check
        PoseEngine(proto, cm, net_size=(100, 52), frame_size=(200, 104), batch=1)
    with pytest.raises(Opb200Error, match="bad batch / lanes / capacity"):
        PoseEngine(proto, cm, net_size=(160, 96), frame_size=(320, 180), batch=1, lanes=0)
    with pytest.raises(Opb200Error, match="Prototxt file not found"):
        PoseEngine("/nonexistent.prototxt", cm, net_size=(160, 96), frame_size=(320, 180), batch=1)
