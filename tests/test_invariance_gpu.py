"""GPU, BASELINE sizes, no oracle needed: frames are independent, so permuting the frames of a batch must permute the
outputs and change no bit -- for the net (all four graphs at their full input sizes, batches large enough for the
frame-interleaved layout and for CTA pairs) and for resize -> nms -> connector on planted maps."""
import numpy as np
import pytest
import torch

import bench as B
from openpose_b200 import modelgen, post, synth
from openpose_b200.net import NetB200

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("model,hw,n", [("body_25", (368, 656), 12), ("coco", (368, 656), 10), ("hand", (368, 368), 16),
                                        ("face", (368, 368), 9)])
def test_net_output_follows_a_permutation_of_the_batch(cuda, model, hw, n):
    proto, cm, _, _ = modelgen.build_files(model, 1234)
    x = torch.from_numpy(synth.frames_to_net_input(synth.frames(n, hw[0], hw[1], 31))).to(cuda)
    perm = torch.from_numpy(np.random.default_rng(1).permutation(n)).to(cuda)
    net = NetB200(proto, cm)
    net.initializationOnThread()
    net.forwardPass(x)
    a = net.getOutputBlobArray().clone()
    net.forwardPass(x[perm].contiguous())
    b = net.getOutputBlobArray().clone()
    net.forwardPass(x[:1].contiguous())                         # and frame 0 alone
    c = net.getOutputBlobArray().clone()
    net.close()
    assert torch.isfinite(a).all()
    assert torch.equal(b, a[perm])
    assert torch.equal(c[0], a[0])


def test_post_processing_follows_a_permutation_of_the_batch(cuda):
    n = 12
    outs = np.concatenate([B.planted_net_outputs(3, p, 40 + p) for p in (1, 8, 20, 45)])      # people per frame vary
    perm = np.random.default_rng(2).permutation(n)
    res = []
    for order in (np.arange(n), perm):
        pp = B.PostPath(n, cuda, capacity=128)
        kp, sc, cnt = pp.run(torch.from_numpy(np.ascontiguousarray(outs[order])).to(cuda))
        res.append((kp.cpu().numpy(), sc.cpu().numpy(), cnt.cpu().numpy(), pp.peaks.cpu().numpy()))
    (kp0, sc0, cnt0, pk0), (kp1, sc1, cnt1, pk1) = res
    assert cnt0.min() >= 1 and np.array_equal(cnt1, cnt0[perm])
    for i in range(n):
        j = perm[i]
        k = int(cnt0[j])
        assert np.array_equal(kp1[i, :k], kp0[j, :k]) and np.array_equal(sc1[i, :k], sc0[j, :k])
        for part in range(25):
            m = int(pk0[j, part, 0, 0])
            assert np.array_equal(pk1[i, part, :m + 1], pk0[j, part, :m + 1])


@pytest.mark.parametrize("model,hw", [("body_25", (368, 656)), ("coco", (184, 328)), ("hand", (368, 368))])
def test_net_output_is_the_same_in_every_batch_size(cuda, model, hw):
    """the small-batch planner changes tiles per work item, Cout tile width and CTA pairing with the batch; none of that
    may change a bit of a frame's output"""
    proto, cm, _, _ = modelgen.build_files(model, 1234)
    x = torch.from_numpy(synth.frames_to_net_input(synth.frames(24, hw[0], hw[1], 33))).to(cuda)
    net = NetB200(proto, cm)
    net.initializationOnThread()
    net.forwardPass(x)
    full = net.getOutputBlobArray().clone()
    for n in (1, 2, 3, 4, 6, 9, 16):
        net.forwardPass(x[:n].contiguous())
        assert torch.equal(net.getOutputBlobArray(), full[:n]), "batch %d" % n
    net.close()
