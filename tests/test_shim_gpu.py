"""GPU: the reference-side C++ bindings (integration/: op::NetB200, Caffe-free op::ArrayCpuGpu, op::*Gpu free
functions), compiled against the reference's OWN headers (oracle/Makefile `shim`), run the hot path end to end
and agree bit-for-bit with the Python mirror of the same C ABI, and with the CPU oracle within tolerance."""
import os
import subprocess

import numpy as np
import pytest
import torch

from openpose_b200 import post, synth
from openpose_b200.net import NetB200
from oracle import ref_net

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "shim_test")


def test_cpp_bindings_run_the_hot_path(tmp_path, cuda):
    if not os.path.exists(BIN):
        pytest.skip("oracle/_ref/shim_test not built (needs /root/reference at build time)")
    n, h, w = 2, 112, 160
    proto, cm, ref = ref_net.build("body_25", 1234)
    x = synth.frames_to_net_input(synth.frames(n, h, w, 9))
    x.tofile(tmp_path / "in.bin")
    prefix = str(tmp_path / "out")
    r = subprocess.run([BIN, proto, cm, str(tmp_path / "in.bin"), str(n), str(h), str(w), prefix],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "net_output %d 78 %d %d" % (n, h // 8, w // 8) in r.stdout
    net_cpp = np.fromfile(prefix + "_net.bin", np.float32).reshape(n, 78, h // 8, w // 8)
    # same library, same inputs -> identical to the Python mirror
    net = NetB200(proto, cm)
    net.initializationOnThread()
    net.forwardPass(x)
    out = net.getOutputBlobArray()
    assert np.array_equal(net_cpp, out.cpu().numpy())
    want = ref.forward(x, keep=["net_output"])["net_output"]
    assert np.abs(net_cpp - want).max() <= 2e-2
    heat = torch.empty((1, 78, h, w), device=cuda)
    post.resize_and_merge_gpu(heat, [out[:1].contiguous()])
    peaks = torch.zeros((1, 25, 128, 3), device=cuda)
    off = float(np.float32(0.5 / float(np.float32(1.95913))))
    post.nms_gpu(peaks, heat, 0.05, (off, off))
    peaks_cpp = np.fromfile(prefix + "_peaks.bin", np.float32).reshape(25, 128, 3)
    pk = peaks[0].cpu().numpy()
    for c in range(25):
        cnt = int(pk[c, 0, 0])
        assert cnt == int(peaks_cpp[c, 0, 0])
        assert np.array_equal(pk[c, 1:cnt + 1], peaks_cpp[c, 1:cnt + 1])
    params = post.make_connect_params(0, scale_factor=1.95913)
    kp, sc, cnt = post.connect_body_parts_gpu(heat, peaks, params, capacity=1024)
    people = int(cnt[0].item())
    assert "people %d" % people in r.stdout
    if people:
        kp_cpp = np.fromfile(prefix + "_kp.bin", np.float32).reshape(people, 25, 3)
        sc_cpp = np.fromfile(prefix + "_sc.bin", np.float32)
        assert np.array_equal(kp_cpp, kp[0, :people].cpu().numpy())
        assert np.array_equal(sc_cpp, sc[0, :people].cpu().numpy())
    net.close()
