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


def bind_host_to_device(index=None):
    """Binds the calling thread to the CPUs next to CUDA device `index` (default: the current device), so that pinned host
    buffers allocated afterwards (first touch) and the copy/submit threads sit on the GPU's own NUMA node.  On a
    two-socket 8-GPU box a rank whose pinned buffer lives on the other socket copies device->host across the socket
    link.  Returns the CPU list it bound to, or None when NVML / the topology is not available (nothing changes then).
    Call it once per rank before allocating pinned memory (bench.py and the CLI do when OGG_NUMA_BIND != 0)."""
    import os
    try:
        import pynvml
        if index is None:
            index = torch.cuda.current_device()
        pynvml.nvmlInit()
        # NVML enumerates in PCI order; map through the bus id torch reports so that CUDA_VISIBLE_DEVICES is honoured
        props = torch.cuda.get_device_properties(index)
        bus = "%08x:%02x:%02x.0" % (getattr(props, "pci_domain_id", 0), props.pci_bus_id, props.pci_device_id)
        h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = [64 * w + b for w in range(words) for b in range(64) if (int(mask[w]) >> b) & 1]
        allowed = os.sched_getaffinity(0)
        cpus = [c for c in cpus if c in allowed]
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return cpus
    except Exception:
        return None


class host_near_device:
    """`with host_near_device(): buf = pinned allocation` -- runs the body bound to the CPUs of the current device's NUMA
    node (see bind_host_to_device) and restores the thread's affinity afterwards, so that only the placement of the host
    pages changes.  `.bound` tells whether the topology was available.  OGG_NUMA_BIND=0 turns it off."""

    def __init__(self, index=None):
        self.index, self.bound, self._saved = index, False, None

    def __enter__(self):
        import os
        if os.environ.get("OGG_NUMA_BIND", "1") != "0":
            self._saved = os.sched_getaffinity(0)
            self.bound = bind_host_to_device(self.index) is not None
        return self

    def __exit__(self, *exc):
        import os
        if self._saved is not None:
            os.sched_setaffinity(0, self._saved)
        return False
