"""Buffer plumbing: PyTorch owns device (and pinned host) memory, libtcb200 does the work.

PyTorch is used here only for allocation, streams and device selection; no torch op is on the
data path. A missing CUDA device is a hard error (there is no CPU fallback).
"""
import numpy as np

from . import _lib


def _torch():
    import torch

    if not torch.cuda.is_available():
        raise RuntimeError(
            "triton_client_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    return torch


def current_device():
    return _torch().cuda.current_device()


def stream_ptr(device):
    """cudaStream_t (as int) of torch's current stream on `device`."""
    return _torch().cuda.current_stream(device).cuda_stream


def synchronize(device):
    _lib.check(_lib.lib.TcbStreamSync(device, stream_ptr(device)))


def device_buffer(nbytes, device):
    """uint8 CUDA tensor of at least `nbytes` (torch's allocator aligns to 512 bytes)."""
    torch = _torch()
    return torch.empty(max(int(nbytes), 1), dtype=torch.uint8, device="cuda:%d" % device)


def host_view_u8(buf):
    """A flat uint8 numpy view of any bytes-like object or ndarray (no copy when contiguous)."""
    if isinstance(buf, np.ndarray):
        return np.ascontiguousarray(buf).view(np.uint8).reshape(-1)
    return np.frombuffer(buf, dtype=np.uint8)


def to_device_u8(buf, device):
    """H2D copy of a host buffer into a fresh uint8 CUDA tensor (synchronous)."""
    src = host_view_u8(buf)
    dst = device_buffer(src.size, device)
    if src.size:
        _lib.check(_lib.lib.TcbMemcpy(device, dst.data_ptr(), src.ctypes.data, src.size,
                                      stream_ptr(device), 1))
    return dst[: src.size]


def to_host(ptr, nbytes, device):
    """D2H copy of `nbytes` from a device pointer into a new uint8 numpy array (synchronous)."""
    nbytes = int(nbytes)
    if nbytes >= (1 << 20):  # large: a recycled page-locked buffer (no staging, no first-touch page faults)
        from .utils.cuda_shared_memory import _host_pool

        out, host_ptr = _host_pool.leased_array(nbytes, np.uint8, [nbytes], min_bytes=1 << 20)
        if out is not None:
            _lib.check(_lib.lib.TcbMemcpy(device, host_ptr, ptr, nbytes, stream_ptr(device), 1))
            return out
    out = np.empty(nbytes, dtype=np.uint8)
    if nbytes:
        _lib.check(_lib.lib.TcbMemcpy(device, out.ctypes.data, ptr, nbytes,
                                      stream_ptr(device), 1))
    return out
