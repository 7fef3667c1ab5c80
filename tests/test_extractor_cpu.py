"""CPU tests of the extractor mirrors' host logic (no GPU): reshape rules of the layer classes, scale-factor
derivations, heat-map range remaps, crop / back-projection math and the reference's error messages."""
import numpy as np
import pytest
import torch

from openpose_b200 import crop_extractors as ce
from openpose_b200 import pose_extractor as pe
from openpose_b200._lib import Opb200Error, POSE_BODY_25


def test_resize_and_merge_top_shape():
    """resizeAndMergeCaffe.cpp:75-82: H' = round((H*netFactor - 1)*scaleFactor) + 1; N merged to 1 on request."""
    assert pe.resize_and_merge_top_shape([1, 78, 46, 82], 8.0, 1.0, True) == [1, 78, 368, 656]
    assert pe.resize_and_merge_top_shape([4, 22, 46, 46], 8.0, 1.0, False) == [4, 22, 368, 368]
    assert pe.resize_and_merge_top_shape([3, 78, 46, 82], 8.0, 1.0, True)[0] == 1
    # the comment's own example: 100 -> 200 pixels is scale 199/99; here netFactor 2: (100*2 - 1)*1 + 1 = 200
    assert pe.resize_and_merge_top_shape([1, 1, 100, 101], 2.0, 1.0, True)[2:] == [200, 202]
    # scaleFactor != 1 (1 / scaleInputToNetInput): round((46*8-1)*1.9592) + 1
    want = int(np.floor(np.float32(367.0) * np.float32(1.9592) + 0.5)) + 1
    assert pe.resize_and_merge_top_shape([1, 78, 46, 82], 8.0, 1.9592, True)[2] == want


def test_nms_and_maximum_top_shapes():
    assert pe.nms_top_shape([1, 78, 368, 656], 127, 25) == [1, 25, 128, 3]      # nmsCaffe.cpp:105-154
    assert pe.nms_top_shape([2, 19, 10, 10], 127, -1) == [2, 18, 128, 3]
    assert pe.maximum_top_shape([5, 22, 368, 368]) == [5, 1, 21, 3]             # maximumCaffe.cpp:60-75


def test_scale_net_to_output_config1():
    """1280x720 frame, net 656x368 (BASELINE config 0): poseExtractorCaffe.cpp:306-310."""
    s = pe.resize_get_scale_factor((1280, 720), (656, 368))
    assert abs(s - min(655 / 1279, 367 / 719)) < 1e-15
    net = (pe.positive_int_round(s * 1280), pe.positive_int_round(s * 720))
    assert net == (653, 368)
    assert abs(pe.resize_get_scale_factor(net, (1280, 720)) - min(1279 / 652, 719 / 367)) < 1e-15


@pytest.mark.parametrize("mode", list(pe.ScaleMode))
def test_heatmaps_export_oracle_matches_scalar_rules(port, mode):
    """poseExtractorNet.cpp:139-227 restated element by element in float32 numpy vs the C oracle."""
    rng = np.random.default_rng(int(mode))
    parts = rng.uniform(-0.3, 1.3, (3, 5, 7)).astype(np.float32)
    bkg = rng.uniform(-0.3, 1.3, (1, 5, 7)).astype(np.float32)
    pafs = rng.uniform(-1.4, 1.4, (4, 5, 7)).astype(np.float32)
    heat = np.concatenate([parts, bkg, pafs])
    got_all = port.heatmaps_export(heat, 3, True, 4, 7, int(mode))
    got = [torch.from_numpy(got_all[:3]), torch.from_numpy(got_all[3:4]), torch.from_numpy(got_all[4:])]
    assert np.array_equal(port.heatmaps_export(heat, 3, True, 4, 4 | 1, int(mode)), np.concatenate([got_all[:3], got_all[4:]]))
    assert port.heatmaps_export(np.concatenate([parts, pafs]), 3, False, 4, 2, int(mode)) is None   # no bkg channel
    f = np.float32
    trunc = lambda v, lo=0.0: np.minimum(np.maximum(v, f(lo)), f(1.0))   # noqa: E731
    def body(v):
        if mode == pe.ScaleMode.NoScale:
            return v
        if mode in (pe.ScaleMode.PlusMinusOne, pe.ScaleMode.PlusMinusOneFixedAspect):
            return trunc(v) * f(2.0) - f(1.0)
        if mode == pe.ScaleMode.UnsignedChar:
            return np.trunc(trunc(v) * f(255.0) + f(0.5)).astype(np.float32)
        return trunc(v)
    def paf(v):
        if mode == pe.ScaleMode.NoScale:
            return v
        if mode in (pe.ScaleMode.ZeroToOne, pe.ScaleMode.ZeroToOneFixedAspect):
            return trunc(v, -1.0) * f(0.5) + f(0.5)
        if mode == pe.ScaleMode.UnsignedChar:
            return np.trunc((trunc(v, -1.0) * f(128.5) + f(128.5)) + f(0.5)).astype(np.float32)
        return trunc(v, -1.0)
    assert np.array_equal(got[0].numpy(), body(parts))
    assert np.array_equal(got[1].numpy(), body(bkg))
    assert np.array_equal(got[2].numpy(), paf(pafs))


def test_pose_extractor_sanity_errors():
    ex = pe.PoseExtractorB200(POSE_BODY_25, "unused.prototxt", "unused.caffemodel", enableNet=False)
    with pytest.raises(Opb200Error, match="initializationOnThread"):
        ex.forwardPass([np.zeros((1, 3, 8, 8), np.float32)], (8, 8), [1.0], np.zeros((1, 78, 1, 1), np.float32))
    ex.initializationOnThread()
    x = np.zeros((1, 3, 16, 16), np.float32)
    with pytest.raises(Opb200Error, match="Empty inputNetData."):
        ex.forwardPass([], (16, 16), [])
    with pytest.raises(Opb200Error, match="Empty inputNetData."):
        ex.forwardPass([np.zeros((0,), np.float32)], (16, 16), [1.0])
    with pytest.raises(Opb200Error, match="must be same than size"):
        ex.forwardPass([x], (16, 16), [1.0, 0.5])
    with pytest.raises(Opb200Error, match="not empty and you have also explicitly chosen"):
        e2 = pe.PoseExtractorB200(POSE_BODY_25, "unused.prototxt", "unused.caffemodel", enableNet=True)
        e2._initialized = True
        e2.forwardPass([x], (16, 16), [1.0], np.zeros((1, 78, 2, 2), np.float32))
    # (the reference raises this text when the net is DISABLED and no heat-maps are given: poseExtractorCaffe.cpp:222-232)
    with pytest.raises(Opb200Error, match="cannot be empty if mEnableNet is true"):
        ex.forwardPass([x], (16, 16), [1.0])
    with pytest.raises(Opb200Error, match="Model not implemented"):
        pe.PoseExtractorB200(7, "a", "b")
    assert ex.get(pe.PoseProperty.NMSThreshold) == 0.05 and ex.get(pe.PoseProperty.ConnectMinSubsetCnt) == 3
    ex.set(pe.PoseProperty.NMSThreshold, 0.1)
    ex.increase(pe.PoseProperty.NMSThreshold, 0.05)
    assert abs(ex.get(pe.PoseProperty.NMSThreshold) - 0.15) < 1e-12
    assert ex.getHeatMapsCopy().size == 0 and ex.getCandidatesCopy() == []


def test_crop_math(port):
    rng = np.random.default_rng(3)
    img = rng.integers(0, 256, (120, 160, 3), dtype=np.uint8)
    # identity-scale crop: side 40 from a 40-pixel rectangle copies the pixels (INTER_LINEAR at integer positions)
    m = ce.crop_matrix((30.0, 20.0, 40.0, 40.0), 40, mirror=False)
    crop = port.crop_warp(img, m, (40, 40))
    assert m.tolist() == [[1.0, 0.0, 30.0], [0.0, 1.0, 20.0]]
    want = img[20:60, 30:70].transpose(2, 0, 1).astype(np.float32) / 256.0 - 0.5
    assert np.array_equal(crop, want.astype(np.float32))
    # mirrored (left hand): column u of the crop is image column x + width - u
    mm = ce.crop_matrix((30.0, 20.0, 40.0, 40.0), 40, mirror=True)
    crop_m = port.crop_warp(img, mm, (40, 40))
    assert mm.tolist() == [[-1.0, 0.0, 70.0], [0.0, 1.0, 20.0]]
    assert np.array_equal(crop_m[:, :, 1:], want[:, :, ::-1][:, :, :-1].astype(np.float32))
    # back-projection: double-precision affine, float result
    peaks = np.array([[10.0, 5.0, 0.7], [0.5, 39.25, 0.1]], np.float32)
    out = ce.connect_keypoints(peaks, mm)
    assert np.array_equal(out, np.array([[60.0, 25.0, 0.7], [69.5, 59.25, 0.1]], np.float32))
    # scaled crop: rectangle 80 -> side 40, scale 2
    m2 = ce.crop_matrix((10.0, 10.0, 80.0, 80.0), 40, mirror=False)
    assert m2[0, 0] == 2.0 and m2[1, 1] == 2.0 and m2[0, 2] == 10.0
    assert ce.recenter((10.0, 20.0, 40.0, 40.0), 60.0, 60.0) == (0.0, 10.0, 60.0, 60.0)
    kp = np.array([[0, 0, 0.25], [0, 0, 0.5], [0, 0, 0.75]], np.float32)
    assert ce.get_average_score(kp) == np.float32(0.5)
    assert np.array_equal(ce.u_char_to_float(np.array([[[0, 128, 255]]], np.uint8)).ravel(),
                          np.array([-0.5, 0.0, 255 / 256 - 0.5], np.float32))


def test_crop_extractor_errors(tmp_path):
    a, b = str(tmp_path / "a.prototxt"), str(tmp_path / "b.caffemodel")
    open(a, "w").close()
    open(b, "w").close()                             # the constructors check the files exist, as NetCaffe's does
    with pytest.raises(Opb200Error, match="Net input and output size must be equal."):
        ce.HandExtractorB200((368, 368), (184, 184), a, b)
    with pytest.raises(Opb200Error, match="Prototxt file not found"):
        ce.HandExtractorB200((368, 368), (368, 368), "missing.prototxt", b)
    h = ce.HandExtractorB200((368, 368), (368, 368), a, b)
    img = np.zeros((64, 64, 3), np.uint8)
    with pytest.raises(Opb200Error, match="must be squared"):
        h.forwardPass([[(0.0, 0.0, 20.0, 30.0), (0.0, 0.0, 20.0, 20.0)]], img)
    with pytest.raises(Opb200Error, match="Empty cvInputData."):
        h.forwardPass([[(0.0, 0.0, 20.0, 20.0), (0.0, 0.0, 20.0, 20.0)]], np.zeros((0, 0, 3), np.uint8))
    h.forwardPass([], img)                           # no people: empty outputs, net untouched
    assert h.getHandKeypoints()[0].shape == (0, 21, 3)
    h.forwardPass([[(0.0, 0.0, 1.0, 1.0), (0.0, 0.0, 0.0, 0.0)]], img)   # rectangles too small: zeros, net untouched
    assert h.getHandKeypoints()[1].shape == (1, 21, 3) and not h.getHandKeypoints()[1].any()
    f = ce.FaceExtractorB200((368, 368), (368, 368), a, b)
    with pytest.raises(Opb200Error, match="Face rectangle for face keypoint estimation must be squared"):
        f.forwardPass([(0.0, 0.0, 50.0, 60.0)], img)
    f.forwardPass([(0.0, 0.0, 40.0, 40.0)], img)     # minFaceSize must exceed 40
    assert f.getFaceKeypoints().shape == (1, 70, 3) and not f.getFaceKeypoints().any()
