"""Device-buffer plumbing: torch owns HBM allocations and streams, nothing else (no torch math)."""
import ctypes

import numpy as np
import torch

from . import _cabi


def ready():
    _cabi.require_gpu()
    if not torch.cuda.is_available():
        raise _cabi.OggError("torch sees no CUDA device")


def is_dev(a):
    return isinstance(a, torch.Tensor)


def empty(*shape):
    return torch.empty(shape, dtype=torch.float64, device="cuda")


def full(shape, value):
    return torch.full(shape, value, dtype=torch.float64, device="cuda")


def to_dev(a):
    """numpy / scalar / tensor -> contiguous float64 CUDA tensor."""
    if is_dev(a):
        if a.dtype != torch.float64 or not a.is_cuda:
            a = a.to(device="cuda", dtype=torch.float64)
        return a.contiguous()
    arr = np.ascontiguousarray(np.asarray(a, dtype=np.float64))
    return torch.from_numpy(arr).to("cuda")


def to_host(t):
    return t.detach().cpu().numpy()


def _check(t):
    # the C ABI takes raw pointers + logical shapes: a strided view would be read / written with the wrong strides
    if not t.is_contiguous():
        raise _cabi.OggError("a non-contiguous tensor reached the C ABI; pass .contiguous() data")
    if t.dtype not in (torch.float64, torch.int64):
        raise _cabi.OggError("the C ABI takes 8-byte elements, got %s" % t.dtype)


def ptr(t):
    if t is None:
        return None
    _check(t)
    return ctypes.c_void_p(t.data_ptr())


def ptr_off(t, elem_offset):
    _check(t)
    return ctypes.c_void_p(t.data_ptr() + 8 * int(elem_offset))


def stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
