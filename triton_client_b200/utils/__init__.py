"""Drop-in for ``tritonclient.utils`` on the BYTES marshalling path, computed on a B200.

Mirrors the names, argument meaning, return conventions and errors of
/root/reference/src/python/library/tritonclient/utils/__init__.py:

* ``serialize_byte_tensor``   (:193-246)  -> 0-d ``np.object_`` array holding one ``bytes``
* ``deserialize_bytes_tensor`` (:249-276) -> 1-D ``np.object_`` array of ``bytes``
* ``serialized_byte_size``     (:43-68)
* ``np_to_triton_dtype`` / ``triton_to_np_dtype`` (:133-190), ``InferenceServerException``
  (:71-130), ``raise_error`` (:36-40)

The per-element work (length prefix + payload scatter, offset chase + gather) runs in
libtcb200.so's CUDA kernels through the C ABI (``include/tcb200.h``); there is no CPU fallback.
What stays on the host is what no kernel can remove: turning Python ``bytes`` objects into one
contiguous buffer and back. Callers that already hold device-resident
payloads should use ``serialize_to_device`` / ``deserialize_from_device`` and skip that.
"""
import struct

import numpy as np

from .. import _lib
from .._device import current_device, device_buffer, stream_ptr, to_device_u8, to_host

_DTYPE_TABLE = (
    (bool, "BOOL"), (np.int8, "INT8"), (np.int16, "INT16"), (np.int32, "INT32"),
    (np.int64, "INT64"), (np.uint8, "UINT8"), (np.uint16, "UINT16"), (np.uint32, "UINT32"),
    (np.uint64, "UINT64"), (np.float16, "FP16"), (np.float32, "FP32"), (np.float64, "FP64"),
)


class InferenceServerException(Exception):
    """Non-success status; same accessors as the reference's exception type."""

    def __init__(self, msg, status=None, debug_details=None):
        self._msg = msg
        self._status = status
        self._debug_details = debug_details

    def __str__(self):
        text = super().__str__() if self._msg is None else self._msg
        return text if self._status is None else "[" + self._status + "] " + text

    def message(self):
        return self._msg

    def status(self):
        return self._status

    def debug_details(self):
        return self._debug_details


def raise_error(msg):
    raise InferenceServerException(msg=msg) from None


def np_to_triton_dtype(np_dtype):
    for np_t, name in _DTYPE_TABLE:
        if np_dtype == np_t:
            return name
    if np_dtype == np.object_ or np_dtype.type == np.bytes_:
        return "BYTES"
    return None


def triton_to_np_dtype(dtype):
    if dtype == "BF16":  # carried as float32 on the client side
        return np.float32
    if dtype == "BYTES":
        return np.object_
    for np_t, name in _DTYPE_TABLE:
        if dtype == name:
            return np_t
    return None


def serialized_byte_size(tensor_value):
    """Total ``len()`` of the items of an ``np.object_`` array (for the 0-d result of
    ``serialize_byte_tensor`` that is the serialized size)."""
    if tensor_value.dtype != np.object_:
        raise_error("The tensor_value dtype must be np.object_")
    if tensor_value.size == 0:
        return 0
    return sum(len(x) for x in tensor_value.ravel(order="C").tolist())


# ---- device-level API ---------------------------------------------------------------------------


def serialize_to_device(payload, lengths, out=None, out_offset=0, device=None, sync=True):
    """Serialise device-resident payloads.

    payload : torch uint8 CUDA tensor, all payloads back to back in element order
    lengths : torch int32/int64 CUDA tensor, one length per element
    out     : None (a new torch uint8 tensor is returned), a torch uint8 CUDA tensor, or a
              ``CudaSharedMemoryRegion``; written from ``out_offset``
    Returns (out, total_bytes). With ``sync=False`` total_bytes is None and nothing waits.
    """
    import torch

    dev = payload.device.index if device is None else device
    n = int(lengths.numel())
    p_bytes = int(payload.numel())
    if out is None:
        out = torch.empty(p_bytes + 4 * n, dtype=torch.uint8, device="cuda:%d" % dev)
    dst_ptr, dst_cap = _dst_of(out, out_offset)
    ws_bytes = int(_lib.lib.TcbBytesSerializeWorkspace(n))
    ws = device_buffer(ws_bytes, dev)
    total = _lib.c_uint64(0)
    lsize = {torch.int32: 4, torch.int64: 8}[lengths.dtype]
    _lib.check(_lib.lib.TcbBytesSerialize(
        dev, payload.data_ptr(), lengths.data_ptr(), lsize, n, dst_ptr, dst_cap, ws.data_ptr(),
        ws_bytes, None, _lib.ctypes.byref(total) if sync else None, stream_ptr(dev)))
    return out, (int(total.value) if sync else None)


def deserialize_from_device(stream, nbytes=None, max_elements=-1, stream_offset=0, device=None):
    """Decode a serialized BYTES stream that lives on the device (a torch uint8 CUDA tensor or a
    ``CudaSharedMemoryRegion``). Returns (offsets int64 CUDA tensor [n+1], payload uint8 CUDA
    tensor): element i is ``payload[offsets[i]:offsets[i+1]]``. Raises ``struct.error`` where the
    reference's ``unpack_from`` would."""
    import torch

    src_ptr, cap = _dst_of(stream, stream_offset)
    dev = _device_of(stream) if device is None else device
    nbytes = cap if nbytes is None else int(nbytes)
    if nbytes > cap:
        raise ValueError("stream of %d bytes exceeds the buffer (%d)" % (nbytes, cap))
    if max_elements is None or max_elements < 0:
        max_elements = -1
    # count first (chase + verification), then allocate exactly: sizing for the worst case of
    # nbytes / 4 elements cost 4.2 GB of int64 offsets for a 2.1 GB stream
    ws_bytes = int(_lib.lib.TcbBytesDeserializeWorkspace(nbytes))
    ws = device_buffer(ws_bytes, dev)
    n = _lib.c_uint64(0)
    bound = _lib.c_uint64(0)
    s = stream_ptr(dev)
    rc = _lib.lib.TcbBytesDeserializeCount(dev, src_ptr, nbytes, int(max_elements), ws.data_ptr(), ws_bytes,
                                           _lib.ctypes.byref(n), _lib.ctypes.byref(bound), s)
    if rc == _lib.TCB_ERR_TRUNCATED:
        raise struct.error(_lib.last_error())
    _lib.check(rc)
    ocap = int(n.value)
    offsets = torch.empty(ocap + 1, dtype=torch.int64, device="cuda:%d" % dev)
    payload = torch.empty(max(int(bound.value), 1), dtype=torch.uint8, device="cuda:%d" % dev)
    pb = _lib.c_uint64(0)
    rc = _lib.lib.TcbBytesDeserializeFill(
        dev, src_ptr, nbytes, int(max_elements), offsets.data_ptr(), ocap, payload.data_ptr(),
        int(payload.numel()), ws.data_ptr(), ws_bytes, _lib.ctypes.byref(n), _lib.ctypes.byref(pb), s)
    if rc == _lib.TCB_ERR_TRUNCATED:
        raise struct.error(_lib.last_error())
    _lib.check(rc)
    return offsets[: int(n.value) + 1], payload[: int(pb.value)]


def _device_of(obj):
    if hasattr(obj, "_device_id"):
        return obj._device_id
    return obj.device.index


def _dst_of(obj, offset):
    """(pointer, capacity) of a torch tensor or a CudaSharedMemoryRegion from ``offset``."""
    if hasattr(obj, "_base_addr"):
        if offset > obj._byte_size:
            raise ValueError("offset beyond the region")
        return int(obj._base_addr) + offset, int(obj._byte_size) - offset
    nbytes = obj.numel() * obj.element_size()
    if offset > nbytes:
        raise ValueError("offset beyond the tensor")
    return obj.data_ptr() + offset, nbytes - offset


# ---- reference-facing API -----------------------------------------------------------------------


from . import _small

try:  # the two per-element loops in C (csrc/pyobj.c, built next to libtcb200.so by `make`)
    from .. import _pyobj
except ImportError:  # not built (no Python headers): the same loops in Python below
    _pyobj = None


def _gather_objects(input_tensor, dev):
    """(device payload tensor, device int64 lengths tensor) of an np.object_ tensor, C order: the
    host half of utils/__init__.py:226-243 of the reference. With _pyobj the elements are visited once
    in C and their bytes copied by a few threads straight into a pooled page-locked buffer."""
    import torch

    n = int(input_tensor.size)
    if _pyobj is None:
        elems = _elements_as_bytes(input_tensor)
        lens = np.fromiter(map(len, elems), dtype=np.int64, count=n)
        if lens.size and int(lens.max()) > 0xFFFFFFFF:
            raise struct.error("'I' format requires 0 <= number <= 4294967295")
        return to_device_u8(b"".join(elems), dev), torch.from_numpy(lens).to("cuda:%d" % dev)
    from .cuda_shared_memory import _host_pool

    flat = np.ascontiguousarray(input_tensor.ravel(order="C"))  # PyObject* items, kept alive until copied
    lens = np.empty(n, dtype=np.int64)
    total, handle = _pyobj.prepare(flat.ctypes.data, n, lens.ctypes.data)
    if n and int(lens.max()) > 0xFFFFFFFF:
        raise struct.error("'I' format requires 0 <= number <= 4294967295")
    d_lens = torch.from_numpy(lens).to("cuda:%d" % dev)
    d_blob = torch.empty(max(total, 1), dtype=torch.uint8, device="cuda:%d" % dev)
    if total:
        stage, ptr = _host_pool.leased_array(total, np.uint8, [total], min_bytes=1 << 20)
        if stage is None:
            stage = np.empty(total, dtype=np.uint8)
            ptr = stage.ctypes.data
        _pyobj.copy(handle, ptr, total, 6)
        _lib.check(_lib.lib.TcbMemcpy(dev, d_blob.data_ptr(), ptr, total, stream_ptr(dev), 1))
        del stage
    return d_blob, d_lens


def _elements_as_bytes(input_tensor):
    """Host-side normalisation of an object array (utils/__init__.py:229-237): exact ``bytes``
    pass through, everything else is ``str()``-encoded as UTF-8; logical C order."""
    out = input_tensor.ravel(order="C").tolist()
    for i, s in enumerate(out):
        if type(s) != bytes:
            out[i] = str(s).encode("utf-8")
    return out


def serialize_byte_tensor_to(out, input_tensor, out_offset=0, device=None):
    """Serialise ``input_tensor`` straight into ``out`` (a ``CudaSharedMemoryRegion`` or torch
    CUDA uint8 tensor) at ``out_offset``; returns the number of bytes written. This is the path
    north_star describes: nothing serialized ever exists on the host."""
    import torch

    if input_tensor.size == 0:
        return 0
    if (input_tensor.dtype != np.object_) and (input_tensor.dtype.type != np.bytes_):
        raise_error("cannot serialize bytes tensor: invalid datatype")
    dev = (_device_of(out) if out is not None else current_device()) if device is None else device
    n = int(input_tensor.size)
    if input_tensor.dtype == np.object_:
        d_blob, d_lens = _gather_objects(input_tensor, dev)
        if out is None:
            return serialize_to_device(d_blob, d_lens, None, 0, dev)
        _, total = serialize_to_device(d_blob, d_lens, out, out_offset, dev)
        return total
    # np.bytes_ (S<k>): a dense n x k byte matrix; lengths (trailing NULs stripped) on the device
    rows = np.ascontiguousarray(input_tensor)
    width = rows.dtype.itemsize
    d_rows = to_device_u8(rows.view(np.uint8).reshape(-1), dev)
    if out is None:
        out_t = torch.empty(n * (width + 4), dtype=torch.uint8, device="cuda:%d" % dev)
    else:
        out_t = out
    dst_ptr, dst_cap = _dst_of(out_t, out_offset if out is not None else 0)
    ws_bytes = int(_lib.lib.TcbBytesSerializeWorkspace(n))
    ws = device_buffer(ws_bytes, dev)
    total = _lib.c_uint64(0)
    _lib.check(_lib.lib.TcbBytesSerializeFixedWidth(
        dev, d_rows.data_ptr(), width, n, dst_ptr, dst_cap, ws.data_ptr(), ws_bytes, None,
        _lib.ctypes.byref(total), stream_ptr(dev)))
    if out is None:
        return out_t, int(total.value)
    return int(total.value)


def serialize_byte_tensor(input_tensor):
    """Same contract as the reference: returns a 0-d ``np.object_`` array whose ``.item()`` is the
    serialized ``bytes`` (empty input: ``np.empty([0], dtype=np.object_)``)."""
    if input_tensor.size == 0:
        return np.empty([0], dtype=np.object_)
    if (input_tensor.dtype != np.object_) and (input_tensor.dtype.type != np.bytes_):
        raise_error("cannot serialize bytes tensor: invalid datatype")
    if _pyobj is not None and input_tensor.dtype == np.object_ and input_tensor.size <= _small.MAX_ELEMENTS:
        # a handful of strings: the kernel works straight on page-locked host memory (utils/_small.py)
        flat = np.ascontiguousarray(input_tensor.ravel(order="C"))
        flattened = _small.serialize_objects(_pyobj, flat, current_device())
        if flattened is not None:
            return np.asarray(flattened, dtype=np.object_)
    out_t, total = serialize_byte_tensor_to(None, input_tensor)
    flattened = to_host(out_t.data_ptr(), total, out_t.device.index).tobytes()
    return np.asarray(flattened, dtype=np.object_)


def _object_array_from(offsets, payload):
    """1-D object array of ``bytes`` from host offsets / payload (the unavoidable per-element
    Python object creation)."""
    n = int(offsets.size) - 1
    arr = np.empty(n, dtype=np.object_)
    if _pyobj is not None:
        off = np.ascontiguousarray(offsets, dtype=np.int64)
        pay = np.ascontiguousarray(payload, dtype=np.uint8)
        _pyobj.scatter(pay.ctypes.data, off.ctypes.data, n, arr.ctypes.data)
        return arr
    blob = payload.tobytes()
    off = offsets.tolist()
    arr[:] = [blob[off[i]:off[i + 1]] for i in range(n)]
    return arr


def deserialize_bytes_tensor(encoded_tensor):
    """Same contract as the reference: any bytes-like object in, 1-D ``np.object_`` array of
    ``bytes`` out; ``struct.error`` on a truncated buffer."""
    buf = memoryview(encoded_tensor).cast("B") if not isinstance(encoded_tensor, (bytes, bytearray)) \
        else encoded_tensor
    if len(buf) == 0:
        return np.array([], dtype=np.object_)
    dev = current_device()
    small = _small.deserialize_stream(buf, dev)
    if small is not None:
        return _object_array_from(small[0], small[1])
    d_stream = to_device_u8(buf, dev)
    offsets, payload = deserialize_from_device(d_stream, len(buf), -1, 0, dev)
    h_off = to_host(offsets.data_ptr(), offsets.numel() * 8, dev).view(np.int64)
    h_pay = to_host(payload.data_ptr(), payload.numel(), dev)
    return _object_array_from(h_off, h_pay)


# ---- BF16 (utils/__init__.py:279-348 of the reference) ---------------------------------------


def serialize_bf16_to_device(values, out=None, out_offset=0, device=None, sync=True):
    """Pack a device-resident float32 tensor into bfloat16 wire bytes (upper 16 bits of every
    float32 pattern, C order). ``values``: contiguous torch float32 CUDA tensor; ``out``: None (a new
    torch uint8 tensor), a torch CUDA tensor or a ``CudaSharedMemoryRegion``, written from
    ``out_offset``. Returns (out, bytes_written)."""
    import torch

    if values.dtype != torch.float32:
        raise_error("cannot serialize bf16 tensor: invalid datatype")
    if not values.is_contiguous():
        raise ValueError("serialize_bf16_to_device needs a C-contiguous tensor")
    dev = values.device.index if device is None else device
    n = int(values.numel())
    if out is None:
        out = torch.empty(2 * n, dtype=torch.uint8, device="cuda:%d" % dev)
    dst_ptr, dst_cap = _dst_of(out, out_offset)
    _lib.check(_lib.lib.TcbBf16Serialize(dev, values.data_ptr(), n, dst_ptr, dst_cap, stream_ptr(dev)))
    if sync:
        _lib.check(_lib.lib.TcbStreamSync(dev, stream_ptr(dev)))
    return out, 2 * n


def deserialize_bf16_from_device(stream, nbytes=None, stream_offset=0, out=None, device=None, sync=True):
    """Expand bfloat16 wire bytes that live on the device (torch uint8 CUDA tensor or a
    ``CudaSharedMemoryRegion``) into a torch float32 CUDA tensor of ``nbytes // 2`` values."""
    import torch

    src_ptr, cap = _dst_of(stream, stream_offset)
    dev = _device_of(stream) if device is None else device
    nbytes = cap if nbytes is None else int(nbytes)
    if nbytes > cap:
        raise ValueError("stream of %d bytes exceeds the buffer (%d)" % (nbytes, cap))
    if out is None:
        out = torch.empty(nbytes // 2, dtype=torch.float32, device="cuda:%d" % dev)
    rc = _lib.lib.TcbBf16Deserialize(dev, src_ptr, nbytes, out.data_ptr(), out.numel() * 4, stream_ptr(dev))
    if rc == _lib.TCB_ERR_TRUNCATED:
        raise struct.error(_lib.last_error())
    _lib.check(rc)
    if sync:
        _lib.check(_lib.lib.TcbStreamSync(dev, stream_ptr(dev)))
    return out


def serialize_bf16_tensor(input_tensor):
    """Same contract as the reference (utils/__init__.py:279-323): a float32 ndarray in, a 0-d
    ``np.object_`` array holding the 2-bytes-per-element wire ``bytes`` out (empty input:
    ``np.empty([0], dtype=np.object_)``). Logical C order, so views and Fortran arrays work."""
    import torch

    if input_tensor.size == 0:
        return np.empty([0], dtype=np.object_)
    if input_tensor.dtype != np.float32:
        raise_error("cannot serialize bf16 tensor: invalid datatype")
    dev = current_device()
    small = _small.bf16_serialize(np.ascontiguousarray(input_tensor), dev)
    if small is not None:
        return np.asarray(small, dtype=np.object_)
    d_src = to_device_u8(np.ascontiguousarray(input_tensor), dev)
    n = int(input_tensor.size)
    out = torch.empty(2 * n, dtype=torch.uint8, device="cuda:%d" % dev)
    _lib.check(_lib.lib.TcbBf16Serialize(dev, d_src.data_ptr(), n, out.data_ptr(), 2 * n, stream_ptr(dev)))
    flattened = to_host(out.data_ptr(), 2 * n, dev).tobytes()  # synchronises
    return np.asarray(flattened, dtype=np.object_)


def deserialize_bf16_tensor(encoded_tensor):
    """Same contract as the reference (utils/__init__.py:326-348), including its result shape:
    the reference appends 1-tuples from ``struct.unpack``, so n values come back as a float32
    array shaped (n, 1) (an empty buffer as shape (0,)); an odd trailing byte raises
    ``struct.error``."""
    buf = memoryview(encoded_tensor).cast("B") if not isinstance(encoded_tensor, (bytes, bytearray)) \
        else encoded_tensor
    if len(buf) == 0:
        return np.array([], dtype=np.float32)
    if len(buf) & 1:
        raise struct.error("unpack_from requires a buffer of at least %d bytes for unpacking 2 bytes at offset %d "
                           "(actual buffer size is %d)" % (len(buf) + 1, len(buf) - 1, len(buf)))
    dev = current_device()
    small = _small.bf16_deserialize(buf, dev)
    if small is not None:
        return small.reshape(-1, 1)
    d_stream = to_device_u8(buf, dev)
    out = deserialize_bf16_from_device(d_stream, len(buf), 0, None, dev, sync=False)
    host = to_host(out.data_ptr(), out.numel() * 4, dev).view(np.float32)
    return host.reshape(-1, 1)
