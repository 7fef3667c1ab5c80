"""Python mirror of the reference's op::Net seam for the B200 backend (C ABI: include/opb200.h, opb200_net_*).

  NetB200(caffeProto, caffeTrainedModel, gpuId=0, enableGoogleLogging=True, lastBlobName="net_output")
      <- op::NetCaffe's constructor signature (reference: include/openpose/net/netCaffe.hpp:12-13)
  .initializationOnThread()   <- NetCaffe::initializationOnThread (src/openpose/net/netCaffe.cpp:154-210)
  .forwardPass(inputData)     <- NetCaffe::forwardPass            (src/openpose/net/netCaffe.cpp:212-261)
  .getOutputBlobArray()       <- NetCaffe::getOutputBlobArray     (src/openpose/net/netCaffe.cpp:263-278)
Same argument meaning and error behaviour (messages included). torch is used only to own device memory and to
hand the output blob back as a zero-copy view; all compute is in libopb200.so. No CPU fallback exists.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

from . import _lib
from ._lib import Opb200Error


class NetInfo(C.Structure):
    _fields_ = [("conv_layers", C.c_int), ("kernel_launches_per_forward", C.c_int),
                ("flops_per_forward", C.c_double), ("arena_bytes", C.c_double), ("output_channels", C.c_int)]


_lib.register({
    "opb200_net_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_char_p, C.c_char_p, C.c_int, C.c_char_p]),
    "opb200_net_destroy": (None, [C.c_void_p]),
    "opb200_net_describe_shape": (C.c_int, [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_char_p,
                                            C.c_size_t]),
    "opb200_net_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, _lib._I4, C.c_void_p]),
    "opb200_net_output": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), _lib._I4]),
    "opb200_net_input_buffer": (C.c_int, [C.c_void_p, _lib._I4, C.POINTER(C.c_void_p), C.c_void_p]),
    "opb200_net_info": (C.c_int, [C.c_void_p, C.POINTER(NetInfo)]),
    "opb200_net_read_blob": (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.c_size_t, _lib._I4]),
    "opb200_net_describe": (C.c_int, [C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_size_t]),
})


class _DevView:
    """Zero-copy description of a device buffer for torch.as_tensor (CUDA array interface v3)."""

    def __init__(self, ptr: int, shape, typestr="<f4"):
        self.__cuda_array_interface__ = {"shape": tuple(int(s) for s in shape), "typestr": typestr,
                                         "data": (int(ptr), False), "version": 3, "strides": None}


def describe_shape(proto: str, caffemodel: str, n: int, h: int, w: int, last_blob: str = "net_output"):
    """Host-only (no CUDA): the activation levels for an (n, 3, h, w) input as dicts with h, w, convs, max_pad, flat, sep --
    which levels are stored frame-interleaved and with how many zero separator columns."""
    L = _lib.lib()
    buf = C.create_string_buffer(1 << 16)
    _lib.check(L.opb200_net_describe_shape(proto.encode(), caffemodel.encode(), last_blob.encode(), int(n), int(h), int(w),
                                           buf, len(buf)))
    levels = []
    for line in buf.value.decode().splitlines():
        levels.append({k: int(v) for k, v in (t.split("=") for t in line.split()[2:])})
    return levels


def describe(proto: str, caffemodel: str, last_blob: str = "net_output") -> str:
    """Host-only plan dump (no CUDA): one line per kernel launch of a forward pass."""
    L = _lib.lib()
    buf = C.create_string_buffer(1 << 20)
    _lib.check(L.opb200_net_describe(proto.encode(), caffemodel.encode(), last_blob.encode(), buf, len(buf)))
    return buf.value.decode()


class NetB200:
    def __init__(self, caffeProto: str, caffeTrainedModel: str, gpuId: int = 0, enableGoogleLogging: bool = True,
                 lastBlobName: str = "net_output"):
        self._proto, self._model, self._gpu, self._last = caffeProto, caffeTrainedModel, int(gpuId), lastBlobName
        self._h: Optional[C.c_void_p] = None
        # NetCaffe checks the two files in its constructor (netCaffe.cpp:60-64)
        hint = ".\nPossible causes:"
        if not os.path.isfile(caffeProto):
            raise Opb200Error("Prototxt file not found: " + caffeProto + hint)
        if not os.path.isfile(caffeTrainedModel):
            raise Opb200Error("Caffe trained model file not found: " + caffeTrainedModel + hint)

    def initializationOnThread(self) -> None:
        L = _lib.lib()
        h = C.c_void_p()
        _lib.check(L.opb200_net_create(C.byref(h), self._proto.encode(), self._model.encode(), self._gpu,
                                       self._last.encode()))
        self._h = h

    def _handle(self):
        if self._h is None:
            raise Opb200Error("NetB200: initializationOnThread() has not been called")
        return self._h

    def forwardPass(self, inputData) -> None:
        """inputData: (N,3,H,W) float32, numpy / CPU torch tensor (copied H2D inside) or CUDA torch tensor."""
        import torch
        L = _lib.lib()
        if inputData is None or getattr(inputData, "size", 1) == 0 or (hasattr(inputData, "numel") and inputData.numel() == 0):
            raise Opb200Error("The Array inputData cannot be empty.")
        if len(inputData.shape) != 4 or inputData.shape[1] != 3:
            raise Opb200Error("The Array inputData must have 4 dimensions: [batch size, 3 (RGB), height, width].")
        stream = torch.cuda.current_stream(self._gpu).cuda_stream
        if isinstance(inputData, torch.Tensor):
            if inputData.dtype != torch.float32 or not inputData.is_contiguous():
                raise Opb200Error("inputData must be contiguous float32")
            ptr, on_dev = inputData.data_ptr(), int(inputData.is_cuda)
        else:
            inputData = np.ascontiguousarray(inputData, np.float32)
            ptr, on_dev = inputData.ctypes.data, 0
        self._keep = inputData   # the H2D copy is asynchronous for pinned memory
        _lib.check(L.opb200_net_forward(self._handle(), ptr, on_dev, _lib.i4(inputData.shape), stream))

    def getOutputBlobArray(self):
        """(N,C,H/8,W/8) float32 CUDA tensor aliasing the net's output blob (non-owning, like the reference's
        ArrayCpuGpu view: valid until the net is destroyed or reshaped by a later forwardPass)."""
        import torch
        L = _lib.lib()
        ptr = C.c_void_p()
        shape = _lib._I4()
        _lib.check(L.opb200_net_output(self._handle(), C.byref(ptr), shape))
        return torch.as_tensor(_DevView(ptr.value, list(shape)), device=torch.device("cuda", self._gpu))

    def info(self) -> NetInfo:
        L = _lib.lib()
        inf = NetInfo()
        _lib.check(L.opb200_net_info(self._handle(), C.byref(inf)))
        return inf

    def readBlob(self, name: str) -> np.ndarray:
        """Test hook: any intermediate blob as float32 NCHW (synchronises the device)."""
        L = _lib.lib()
        shape = _lib._I4()
        _lib.check(L.opb200_net_read_blob(self._handle(), name.encode(), None, 0, shape))
        out = np.empty(tuple(shape), np.float32)
        _lib.check(L.opb200_net_read_blob(self._handle(), name.encode(), out.ctypes.data, out.size, shape))
        return out

    def close(self) -> None:
        if self._h is not None:
            _lib.lib().opb200_net_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ----------------------------------------------------------------------------------------------------
# bench / smoke helpers (random-init weights: modelgen.py; the oracle is imported by smoke() only, as the checker)
# ----------------------------------------------------------------------------------------------------
class _BenchNet:
    def __init__(self, net: NetB200, batch: int):
        self.net, self.batch = net, batch
        self._out = None

    def forward_device(self, x, ev=None):
        import torch
        if ev is not None:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
        self.net.forwardPass(x)
        if ev is not None:
            b.record()
            ev.append(("net_convs", a, b))
        if self._out is None:
            self._out = self.net.getOutputBlobArray()
        return self._out

    def launches_per_step(self) -> int:
        return self.net.info().kernel_launches_per_forward

    def output_bytes(self) -> int:
        return self._out.numel() * 4 if self._out is not None else 0

    def copy_output_to_host(self):
        import torch
        if getattr(self, "_host", None) is None:
            self._host = torch.empty(self._out.shape, dtype=torch.float32).pin_memory()
        self._host.copy_(self._out, non_blocking=True)


def bench_net(batch: int, h: int, w: int, device, model: str = "body_25", seed: int = 1234) -> _BenchNet:
    from . import modelgen       # seeded random-init caffemodel (there is no network for the trained one)
    proto, cm, _, _ = modelgen.build_files(model, seed)
    net = NetB200(proto, cm, gpuId=device.index or 0)
    net.initializationOnThread()
    return _BenchNet(net, batch)


def smoke() -> None:
    """One small BODY_25 forward on cuda:0 checked against the FP32 CPU oracle (called by __graft_entry__.smoke)."""
    import torch
    from oracle import ref_net
    from . import synth
    proto, cm, ref = ref_net.build("body_25", 1234)
    x = synth.frames_to_net_input(synth.frames(1, 64, 96, 3))
    net = NetB200(proto, cm)
    net.initializationOnThread()
    net.forwardPass(x)
    got = net.getOutputBlobArray().cpu().numpy()
    want = ref.forward(x, keep=["net_output"])["net_output"]
    err = float(np.abs(got - want).max())
    assert got.shape == want.shape and err <= 2e-2, "NetB200 differs from the oracle: max-abs %g" % err
    print("net smoke ok: BODY_25 64x96, %d conv launches, max-abs vs fp32 oracle %.4f" % (net.info().conv_layers, err))
    net.close()
