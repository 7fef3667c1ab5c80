"""GPU: life-cycle of the library's objects -- creating and destroying nets / engines repeatedly gives the device memory
back, and two nets driven from two host threads (one NetB200 per worker thread, as the reference uses NetCaffe) return
what they return when used one after the other."""
import threading

import numpy as np
import pytest
import torch

from openpose_b200 import modelgen, synth
from openpose_b200.crop_extractors import crop_matrix
from openpose_b200.engine import CropEngine, PoseEngine
from openpose_b200.net import NetB200

pytestmark = pytest.mark.gpu


def _free_bytes():
    torch.cuda.synchronize()
    return torch.cuda.mem_get_info()[0]


def test_net_create_destroy_returns_device_memory(cuda):
    proto, cm, _, _ = modelgen.build_files("body_25", 1234)
    x = torch.from_numpy(synth.frames_to_net_input(synth.frames(2, 64, 96, 3))).to(cuda)

    def cycle():
        net = NetB200(proto, cm)
        net.initializationOnThread()
        net.forwardPass(x)
        out = net.getOutputBlobArray().cpu().numpy().copy()
        net.forwardPass(x[:1])                                   # a reshape inside the net's life
        net.close()
        return out

    first = cycle()                                              # (first use also loads modules / creates the context pools)
    base = _free_bytes()
    for _ in range(4):
        assert np.array_equal(cycle(), first)
    assert base - _free_bytes() < (8 << 20), "device memory not returned by opb200_net_destroy"


def test_engine_create_destroy_returns_device_memory(cuda):
    proto, cm, _, _ = modelgen.build_files("body_25", 1234)
    frames = synth.frames(2, 90, 160, 9)

    def cycle():
        eng = PoseEngine(proto, cm, gpus=[0], net_size=(160, 96), frame_size=(160, 90), batch=2, lanes=2, capacity=32)
        eng.submit(0, frames)
        first, kp, sc, cnt = eng.next()
        res = (first, kp.copy(), sc.copy(), cnt.copy())
        eng.close()
        return res

    a = cycle()
    base = _free_bytes()
    for _ in range(3):
        b = cycle()
        assert a[0] == b[0] and all(np.array_equal(u, v) for u, v in zip(a[1:], b[1:]))
    assert base - _free_bytes() < (8 << 20), "device memory not returned by opb200_engine_destroy"


def test_crop_engine_create_destroy_returns_device_memory(cuda):
    proto, cm, _, _ = modelgen.build_files("hand", 4321)
    frame = synth.frames(1, 240, 320, 5)[0]
    mats = np.stack([crop_matrix((40.0, 30.0, 120.0, 120.0), 96, False), crop_matrix((150.0, 60.0, 90.0, 90.0), 96, True)])

    def cycle():
        eng = CropEngine(proto, cm, 22, gpus=[0], frame_size=(320, 240), net_side=96, batch=2, lanes=2)
        eng.submit(0, frame, mats)
        res = eng.next()[1]
        eng.close()
        return res

    a = cycle()
    base = _free_bytes()
    for _ in range(3):
        assert np.array_equal(cycle(), a)
    assert base - _free_bytes() < (8 << 20), "device memory not returned by opb200_engine_destroy (crop mode)"


def test_two_nets_on_two_host_threads(cuda):
    proto, cm, _, _ = modelgen.build_files("body_25", 1234)
    xs = [torch.from_numpy(synth.frames_to_net_input(synth.frames(2, 64, 96, 20 + i))).to(cuda) for i in range(2)]
    want = []
    for x in xs:
        net = NetB200(proto, cm)
        net.initializationOnThread()
        net.forwardPass(x)
        want.append(net.getOutputBlobArray().cpu().numpy().copy())
        net.close()
    got, errors = [None, None], []

    def worker(i):
        try:
            net = NetB200(proto, cm)
            net.initializationOnThread()
            for _ in range(10):
                net.forwardPass(xs[i])
                out = net.getOutputBlobArray().cpu().numpy().copy()
                assert np.array_equal(out, want[i])
            got[i] = out
            net.close()
        except Exception as e:                                   # noqa: BLE001 - reported below
            errors.append(repr(e))

    threads = [threading.Thread(target=worker, args=(i,)) for i in range(2)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    assert all(np.array_equal(g, w) for g, w in zip(got, want))
