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


def ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def ptr_off(t, elem_offset):
    return ctypes.c_void_p(t.data_ptr() + 8 * int(elem_offset))


def stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
