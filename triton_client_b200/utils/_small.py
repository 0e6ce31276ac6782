"""Small BYTES tensors: the same kernels, fed from and writing to page-locked host memory.

A request usually carries a handful of strings. For those the general path - H2D copies of payload
and lengths, the kernel, D2H copies of the result, each synchronised, torch allocations in between -
costs 110 us (serialise) / 170-230 us (deserialise) however small the tensor is. Page-locked host
memory is addressable by the device (unified addressing), so for tensors of up to a few hundred KB
the kernels read their input from and write their output to one cached staging buffer directly over
PCIe: one call into the library and one synchronisation per tensor, no copies to issue, nothing to
allocate. Same entry points (TcbBytesSerialize / TcbBytesDeserialize), same results.
"""
import ctypes
import struct
import threading

import numpy as np

from .. import _lib
from .._device import stream_ptr

MAX_ELEMENTS = 16384
MAX_BYTES = 256 << 10          # payload (serialise) / stream (deserialise) bytes handled here
_STAGE_BYTES = 4 << 20         # >= 8n + payload + (payload + 4n), and >= stream + 8(stream/4 + 1) + stream

_lock = threading.Lock()
_state = {}                    # device -> (pinned torch tensor, numpy view, ser workspace, des workspace)


def _get(dev):
    st = _state.get(dev)
    if st is None:
        import torch

        pin = torch.empty(_STAGE_BYTES, dtype=torch.uint8, pin_memory=True)
        ws_ser = torch.empty(max(int(_lib.lib.TcbBytesSerializeWorkspace(MAX_ELEMENTS)), 1), dtype=torch.uint8,
                             device="cuda:%d" % dev)
        ws_des = torch.empty(max(int(_lib.lib.TcbBytesDeserializeWorkspace(MAX_BYTES)), 1), dtype=torch.uint8,
                             device="cuda:%d" % dev)
        d_stream = torch.empty(MAX_BYTES + 16, dtype=torch.uint8, device="cuda:%d" % dev)
        st = (pin, pin.numpy(), ws_ser, ws_des, d_stream)
        _state[dev] = st
    return st


def serialize_objects(pyobj, flat, dev):
    """Serialized bytes of the C-contiguous 1-D np.object_ array `flat`, or None when the tensor is
    too large for this path. `pyobj` is the _pyobj module."""
    n = int(flat.size)
    if n > MAX_ELEMENTS:
        return None
    with _lock:
        pin, view, ws = _get(dev)[:3]
        base = pin.data_ptr()
        total, handle = pyobj.prepare(flat.ctypes.data, n, base)  # lengths (int64) at the start of the stage
        if total > MAX_BYTES:
            return None
        if n and int(view[:8 * n].view(np.int64).max()) > 0xFFFFFFFF:
            raise struct.error("'I' format requires 0 <= number <= 4294967295")
        a = (8 * n + 15) & ~15
        b = (a + total + 15) & ~15
        out_bytes = total + 4 * n
        pyobj.copy(handle, base + a, total, 1)
        done = ctypes.c_uint64(0)
        _lib.check(_lib.lib.TcbBytesSerialize(dev, base + a, base, 8, n, base + b, out_bytes, ws.data_ptr(),
                                              int(ws.numel()), None, ctypes.byref(done), stream_ptr(dev)))
        assert done.value == out_bytes
        return view[b:b + out_bytes].tobytes()


def deserialize_stream(buf, dev, with_fields=False):
    """(offsets int64 [n + 1], payload uint8) as numpy copies for the serialized buffer `buf`
    (bytes-like, at most MAX_BYTES), or None when it is too large for this path."""
    nbytes = len(buf)
    if nbytes > MAX_BYTES:
        return None
    with _lock:
        pin, view, _, ws, d_stream = _get(dev)
        base = pin.data_ptr()
        view[:nbytes] = np.frombuffer(buf, dtype=np.uint8)
        # the chase follows a chain of dependent loads: over PCIe that is ~1.5 us per element (measured:
        # 256 elements 383 us, 4096 elements 1.66 ms), so the stream itself goes to device memory first
        # (one asynchronous copy on the same stream); offsets and payload are written straight to the host
        s = stream_ptr(dev)
        _lib.check(_lib.lib.TcbMemcpy(dev, d_stream.data_ptr(), base, nbytes, s, 0))
        ocap = nbytes // 4 + 1
        o = (nbytes + 15) & ~15
        p = (o + 8 * (ocap + 1) + 15) & ~15
        n, pb = ctypes.c_uint64(0), ctypes.c_uint64(0)
        rc = _lib.lib.TcbBytesDeserialize(dev, d_stream.data_ptr(), nbytes, -1, base + o, ocap, base + p, nbytes,
                                          ws.data_ptr(), int(ws.numel()), ctypes.byref(n), ctypes.byref(pb), s)
        if rc == _lib.TCB_ERR_TRUNCATED:
            raise struct.error(_lib.last_error())
        _lib.check(rc)
        offsets = view[o:o + 8 * (int(n.value) + 1)].view(np.int64).copy()
        payload = view[p:p + int(pb.value)].copy()
        if not with_fields:
            return offsets, payload
        # classification strings: the ':' split (TcbClassificationFields) on what the decoder just wrote,
        # still in the stage - a warp per element reads it over PCIe, 32 bytes per round
        cnt = int(n.value)
        f = (p + int(pb.value) + 15) & ~15
        if cnt:
            _lib.check(_lib.lib.TcbClassificationFields(dev, base + p, base + o, cnt, base + f, s))
            _lib.check(_lib.lib.TcbStreamSync(dev, s))
        fields = view[f:f + 32 * cnt].view(np.int64).reshape(cnt, 4).copy()
        return offsets, payload, fields


BF16_MAX_ELEMENTS = 256 << 10  # 1 MiB of float32 in, 512 KiB out: both fit the stage


def bf16_serialize(values, dev):
    """bfloat16 wire bytes of the C-contiguous float32 ndarray `values` (any shape), or None when it is
    too large for this path. One launch that reads and writes the page-locked stage, one sync."""
    n = int(values.size)
    if n > BF16_MAX_ELEMENTS:
        return None
    with _lock:
        pin, view = _get(dev)[:2]
        base = pin.data_ptr()
        view[:4 * n] = values.reshape(-1).view(np.uint8)
        o = (4 * n + 15) & ~15
        s = stream_ptr(dev)
        _lib.check(_lib.lib.TcbBf16Serialize(dev, base, n, base + o, 2 * n, s))
        _lib.check(_lib.lib.TcbStreamSync(dev, s))
        return view[o:o + 2 * n].tobytes()


def bf16_deserialize(buf, dev):
    """float32 ndarray [n] for the bfloat16 wire bytes `buf` (even length), or None when too large."""
    nbytes = len(buf)
    if nbytes > 2 * BF16_MAX_ELEMENTS:
        return None
    with _lock:
        pin, view = _get(dev)[:2]
        base = pin.data_ptr()
        view[:nbytes] = np.frombuffer(buf, dtype=np.uint8)
        o = (nbytes + 15) & ~15
        s = stream_ptr(dev)
        rc = _lib.lib.TcbBf16Deserialize(dev, base, nbytes, base + o, 2 * nbytes, s)
        if rc == _lib.TCB_ERR_TRUNCATED:
            raise struct.error(_lib.last_error())
        _lib.check(rc)
        _lib.check(_lib.lib.TcbStreamSync(dev, s))
        return view[o:o + 2 * nbytes].view(np.float32).copy()
