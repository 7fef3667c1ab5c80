"""GPU parity: people post-step kernels (through the C ABI) vs the reference's own object code -- bit-exact.
The committed fixture was produced by oracle/_ref; when that library is present it is also called live."""
import os

import numpy as np
import pytest
import torch

from oracle.post import Ref
from openpose_b200 import people as pp
from tests.cases_people import PEOPLE_CASES, SCALE_CASES, TOP_N, make_people

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "people_golden.npz")
CAP = 64


def _batch(cuda):
    """All BODY_25 cases as one batch (N frames, capacity CAP) + the per-frame host arrays."""
    frames = [c for c in PEOPLE_CASES if c[2] == 25]
    kp = np.full((len(frames), CAP, 25, 3), 7.0, np.float32)          # stale rows beyond count must be ignored
    sc = np.full((len(frames), CAP), 9.0, np.float32)
    cnt = np.zeros((len(frames),), np.int32)
    host = []
    for i, (name, model, parts, people, seed) in enumerate(frames):
        k, s = make_people(parts, people, seed)
        kp[i, :people], sc[i, :people], cnt[i] = k, s, people
        host.append((name, k, s))
    return torch.from_numpy(kp).to(cuda), torch.from_numpy(sc).to(cuda), torch.from_numpy(cnt).to(cuda), host


def test_rectangles_match_reference(cuda):
    g = np.load(GOLDEN)
    kp, sc, cnt, host = _batch(cuda)
    hands, faces = pp.hand_face_rectangles_gpu(kp, cnt, PEOPLE_CASES[0][1])
    for i, (name, k, s) in enumerate(host):
        n = k.shape[0]
        assert np.array_equal(hands[i, :n].cpu().numpy(), g[name + "/hands"])
        assert np.array_equal(faces[i, :n].cpu().numpy(), g[name + "/faces"])
        assert not hands[i, n:].any() and not faces[i, n:].any()
    # COCO_18 part indexes
    name, model, parts, people, seed = PEOPLE_CASES[1]
    k, s = make_people(parts, people, seed)
    kk = torch.from_numpy(k[None]).to(cuda).contiguous()
    c1 = torch.tensor([people], dtype=torch.int32, device=cuda)
    assert np.array_equal(pp.HandDetector(model).detectHands(kk, c1)[0].cpu().numpy(), g[name + "/hands"])
    assert np.array_equal(pp.FaceDetector(model).detectFaces(kk, c1)[0].cpu().numpy(), g[name + "/faces"])


@pytest.mark.parametrize("n_max", TOP_N)
def test_keep_top_people_matches_reference(cuda, n_max):
    g = np.load(GOLDEN)
    kp, sc, cnt, host = _batch(cuda)
    out_kp, out_sc, out_n = pp.KeepTopNPeople(n_max).keepTopPeople(kp, sc, cnt)
    for i, (name, k, s) in enumerate(host):
        want = g["%s/top%d" % (name, n_max)]
        assert int(out_n[i]) == want.shape[0]
        got = out_kp[i, :want.shape[0]].cpu().numpy()
        assert np.array_equal(got, want)
        # the scores travel with their people
        kept = [int(np.flatnonzero((k == row).all(axis=(1, 2)))[0]) for row in got]
        assert np.array_equal(out_sc[i, :want.shape[0]].cpu().numpy(), s[kept])


@pytest.mark.parametrize("i", range(len(SCALE_CASES)))
def test_scale_keypoints_matches_reference(cuda, i):
    g = np.load(GOLDEN)
    mode, s_io, s_no, prod = SCALE_CASES[i]
    kp, sc, cnt, host = _batch(cuda)
    before = kp.clone()
    pp.KeypointScaler(mode).scale(kp, s_io, s_no, prod, count=cnt)
    for f, (name, k, s) in enumerate(host):
        n = k.shape[0]
        assert np.array_equal(kp[f, :n].cpu().numpy(), g["%s/scale%d" % (name, i)])
        assert torch.equal(kp[f, n:], before[f, n:])                     # rows beyond count untouched
    # count = None scales every row (hand / face keypoint arrays)
    full = before.clone()
    pp.KeypointScaler(mode).scale(full, s_io, s_no, prod)
    assert torch.equal(full[0, :host[0][1].shape[0]], kp[0, :host[0][1].shape[0]])


@pytest.mark.skipif(not Ref.available(), reason="reference object code not present")
def test_random_people_against_live_reference(cuda):
    ref = Ref()
    rng = np.random.default_rng(11)
    for trial in range(6):
        people = int(rng.integers(1, 60))
        k, s = make_people(25, people, 100 + trial)
        kp = torch.zeros((1, CAP, 25, 3), device=cuda)
        kp[0, :people] = torch.from_numpy(k).to(cuda)
        sc = torch.zeros((1, CAP), device=cuda)
        sc[0, :people] = torch.from_numpy(s).to(cuda)
        cnt = torch.tensor([people], dtype=torch.int32, device=cuda)
        hands, faces = pp.hand_face_rectangles_gpu(kp, cnt, 0)
        assert np.array_equal(hands[0, :people].cpu().numpy(), ref.detect_hands(k, 0))
        assert np.array_equal(faces[0, :people].cpu().numpy(), ref.detect_faces(k, 0))
        n_max = int(rng.integers(1, 12))
        out_kp, _, out_n = pp.KeepTopNPeople(n_max).keepTopPeople(kp, sc, cnt)
        want = ref.keep_top_people(k, s, n_max)
        assert int(out_n[0]) == want.shape[0] and np.array_equal(out_kp[0, :want.shape[0]].cpu().numpy(), want)
