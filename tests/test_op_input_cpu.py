"""Frame -> net input, host side: ScaleAndSizeExtractor arithmetic and the pre-processing oracle's properties
(reference: src/openpose/core/scaleAndSizeExtractor.cpp, cvMatToOpInput.cpp:108-148, gpu/cuda.cu:11-25,
net/resizeAndMergeBase.cu:56-78)."""
import numpy as np
import pytest

from openpose_b200._lib import Opb200Error
from openpose_b200.op_input import ScaleAndSizeExtractor


def test_extract_default_body25_720p():
    # --net_resolution -1x368 on a 1280x720 frame: 16*round(368*1280/720/16) = 656
    scales, sizes, scale_out, out_res = ScaleAndSizeExtractor((-1, 368)).extract((1280, 720))
    assert sizes == [(656, 368)]
    assert scales == [min(655 / 1279.0, 367 / 719.0)]
    assert scale_out == 1.0 and out_res == (1280, 720)


def test_extract_multi_scale_and_output_size():
    ex = ScaleAndSizeExtractor((656, 368), -1.0, (1920, 1080), 3, 0.25)
    scales, sizes, scale_out, out_res = ex.extract((640, 480))
    # 656*0.75 = 492 -> 480; 368*0.75 = 276 -> 272; 656*0.5 = 328 -> 320; 368*0.5 = 184 -> 176
    assert sizes == [(656, 368), (480, 272), (320, 176)]
    for s, (w, h) in zip(scales, sizes):
        assert s == min((w - 1) / 639.0, (h - 1) / 479.0)
    assert out_res == (1920, 1080) and scale_out == min(1919 / 639.0, 1079 / 479.0)


def test_extract_dynamic_behaviour_limits_width():
    # a 4:1 panorama: -1x368 would give 1472 wide; dynamic behaviour 1 caps it at 368*16/9 -> 656
    assert ScaleAndSizeExtractor((-1, 368), -1.0).extract((4000, 1000))[1] == [(1472, 368)]
    assert ScaleAndSizeExtractor((-1, 368), 1.0).extract((4000, 1000))[1] == [(656, 368)]
    assert ScaleAndSizeExtractor((368, -1), 1.0).extract((1000, 4000))[1] == [(368, 208)]


@pytest.mark.parametrize("args,inp,text", [
    (((650, 368),), None, "multiples of 16"),
    (((656, 368), -1.0, (-1, -1), 0), None, "at least 1 scale"),
    (((656, 368), -1.0, (-1, -1), 1, 0.0), None, "strictly positive"),
    (((-1, -1),), (640, 480), "Only 1 of the dimensions"),
    (((656, 368),), (0, 480), "empty cvInputData"),
    (((656, 368), -1.0, (-1, -1), 6, 0.25), (640, 480), "range [0, 1]"),
])
def test_extract_errors(args, inp, text):
    with pytest.raises(Opb200Error) as e:
        ex = ScaleAndSizeExtractor(*args)
        ex.extract(inp)
    assert text in str(e.value)


def test_oracle_scale_one_is_the_normalised_frame(port):
    rng = np.random.default_rng(0)
    frame = rng.integers(0, 256, size=(20, 31, 3), dtype=np.uint8)
    out = port.preprocess(frame, (24, 40), 1.0)
    want = frame.astype(np.float32).transpose(2, 0, 1) / 256.0 - 0.5
    assert np.array_equal(out[:, :20, :31], want)          # dx = dy = 0: the centre tap, exactly
    assert not out[:, 20:, :].any() and not out[:, :, 31:].any()  # padding


def test_oracle_constant_frame_and_padding_edge(port):
    frame = np.full((48, 64, 3), 200, np.uint8)
    frame[..., 1] = 17
    scale = 0.6
    out = port.preprocess(frame, (32, 48), scale)
    vals = np.float32([200, 17, 200]) / np.float32(256) - np.float32(0.5)
    inside_w = int(np.ceil(np.float32(64) * np.float32(scale)))   # x < 38.4 -> 39 columns
    inside_h = int(np.ceil(np.float32(48) * np.float32(scale)))   # y < 28.8 -> 29 rows
    assert inside_w == 39 and inside_h == 29
    for c in range(3):
        assert np.allclose(out[c, :inside_h, :inside_w], vals[c], atol=1e-6)
        assert not out[c, inside_h:, :].any() and not out[c, :, inside_w:].any()


def test_oracle_matches_numpy_restatement(port):
    """Independent float32 numpy evaluation of the same two kernels on a small frame."""
    rng = np.random.default_rng(3)
    hs, ws, ht, wt, scale = 9, 13, 16, 16, np.float32(1.2307693)
    frame = rng.integers(0, 256, size=(hs, ws, 3), dtype=np.uint8)
    planar = frame.astype(np.float32).transpose(2, 0, 1) * np.float32(1 / 256.0) - np.float32(0.5)
    f = np.float32

    def cubic(v0, v1, v2, v3, d):
        return ((f(-0.5) * v0 + f(1.5) * v1 - f(1.5) * v2 + f(0.5) * v3) * d * d * d
                + (v0 - f(2.5) * v1 + f(2) * v2 - f(0.5) * v3) * d * d - f(0.5) * (v0 - v2) * d + v1)

    want = np.zeros((3, ht, wt), np.float32)
    for y in range(ht):
        for x in range(wt):
            if not (f(x) < f(ws) * scale and f(y) < f(hs) * scale):
                continue
            xs = (f(x) + f(0.5)) / scale - f(0.5)
            ys = (f(y) + f(0.5)) / scale - f(0.5)
            x1 = min(ws - 1, max(0, int(np.floor(xs))))
            y1 = min(hs - 1, max(0, int(np.floor(ys))))
            xi = [max(0, x1 - 1), x1, min(ws - 1, x1 + 1), min(ws - 1, min(ws - 1, x1 + 1) + 1)]
            yi = [max(0, y1 - 1), y1, min(hs - 1, y1 + 1), min(hs - 1, min(hs - 1, y1 + 1) + 1)]
            dx, dy = xs - f(x1), ys - f(y1)
            for c in range(3):
                t = [cubic(*[planar[c, yy, xx] for xx in xi], dx) for yy in yi]
                want[c, y, x] = cubic(*t, dy)
    assert np.array_equal(port.preprocess(frame, (ht, wt), float(scale), contract=False), want)
