"""Drop-in for ``tritonclient.utils.shared_memory`` (system / POSIX shared memory), SURVEY.md 8f row f2.

Same functions, argument order, defaults and exception texts as
/root/reference/src/python/library/tritonclient/utils/shared_memory/__init__.py (:93 create, :129
set, :186 get_contents_as_numpy, :262 mapped_shared_memory_regions, :274 destroy, :314
SharedMemoryException); the C side is libtcb200.so's ``TcbShmRegion*`` instead of libcshm.

Additions (optional arguments / new functions, existing call sites keep working):
* the mapping is page-locked when a CUDA device is present, and ``set_shared_memory_region_from_device``
  / ``get_contents_to_device`` move tensors between the device and the region by DMA;
* BYTES contents (``np.object_``) are decoded by the CUDA kernels when a device is present (the
  reference loops over the elements in Python), with the reference's ``prod(shape)`` quirks kept;
* copies are bounds-checked.
"""
import ctypes
import struct

import numpy as np

from ... import _lib

mapped_shm_regions = []


class SharedMemoryException(Exception):
    """Exception indicating non-Success status (reference :314-340: code -> fixed text)."""

    def __init__(self, err):
        self.err_code_map = {
            -2: "unable to get shared memory descriptor",
            -3: "unable to initialize the size",
            -4: "unable to read/mmap the shared memory region",
            -5: "unable to unlink the shared memory region",
            -6: "unable to munmap the shared memory region",
        }
        self._msg = None
        if type(err) == str:
            self._msg = err
        else:
            code = err.value if hasattr(err, "value") else int(err)
            if code != 0 and code in self.err_code_map:
                self._msg = self.err_code_map[code]

    def __str__(self):
        return super().__str__() if self._msg is None else self._msg


def _raise_if_error(rc, libcshm_codes=False):
    """libcshm's own failure codes (create / destroy) map to the reference's fixed texts; everything
    else (bounds checks, CUDA errors) carries the library's message."""
    if rc != 0:
        if libcshm_codes and rc in (-2, -3, -4, -5, -6):
            raise SharedMemoryException(ctypes.c_int(rc))
        raise SharedMemoryException(_lib.last_error())


def _raise_error(msg):
    raise SharedMemoryException(msg)


def _info(shm_handle):
    addr, key, fd = ctypes.c_void_p(), ctypes.c_char_p(), ctypes.c_int()
    off, size, pinned = ctypes.c_uint64(), ctypes.c_uint64(), ctypes.c_int()
    _raise_if_error(_lib.lib.TcbShmRegionInfo(shm_handle, ctypes.byref(addr), ctypes.byref(key), ctypes.byref(fd),
                                              ctypes.byref(off), ctypes.byref(size), ctypes.byref(pinned)))
    return addr.value, key.value.decode("utf-8"), off.value, size.value, bool(pinned.value)


def _has_device():
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


def create_shared_memory_region(triton_shm_name, shm_key, byte_size, pin=None):
    """shm_open(shm_key) + ftruncate + mmap; returns the handle (a ``c_void_p``). ``pin`` (default:
    when a CUDA device is present) page-locks the mapping for DMA."""
    shm_handle = ctypes.c_void_p()
    if pin is None:
        pin = _has_device()
    _raise_if_error(_lib.lib.TcbShmRegionCreate(triton_shm_name.encode("utf-8"), shm_key.encode("utf-8"),
                                                int(byte_size), 1 if pin else 0, ctypes.byref(shm_handle)), libcshm_codes=True)
    mapped_shm_regions.append(shm_key)
    return shm_handle


def set_shared_memory_region(shm_handle, input_values, offset=0):
    """Copy the numpy arrays back to back into the region from ``offset``; an ``np.object_`` array is
    expected to be the result of ``serialize_byte_tensor``."""
    if not isinstance(input_values, (list, tuple)):
        _raise_error("input_values must be specified as a list/tuple of numpy arrays")
    for input_value in input_values:
        if not isinstance(input_value, np.ndarray):
            _raise_error("each element of input_values must be a numpy array")
    offset_current = offset
    for input_value in input_values:
        input_value = np.ascontiguousarray(input_value).reshape(-1)
        if input_value.dtype == np.object_:
            data = input_value.item()
            byte_size = len(data)
            _raise_if_error(_lib.lib.TcbShmRegionSet(shm_handle, offset_current, byte_size,
                                                     ctypes.cast(ctypes.c_char_p(data), ctypes.c_void_p)))
        else:
            byte_size = input_value.size * input_value.itemsize
            _raise_if_error(_lib.lib.TcbShmRegionSet(shm_handle, offset_current, byte_size,
                                                     input_value.ctypes.data_as(ctypes.c_void_p)))
        offset_current += byte_size
    return


def get_contents_as_numpy(shm_handle, datatype, shape, offset=0):
    """numpy array over the region's contents (reference :186-259). Fixed-size dtypes: a zero-copy
    view of the mapping, like the reference. BYTES: exactly ``prod(shape)`` elements are chased from
    ``offset`` - never fewer than one, the reference's loop quirk - on the GPU when there is one
    (the mapping is page-locked, the kernels read it over PCIe once), else with the reference's loop."""
    addr, _, region_offset, byte_size, pinned = _info(shm_handle)
    start_pos = region_offset + offset
    if (datatype != np.object_) and (datatype != np.bytes_):
        requested_byte_size = int(np.prod(shape)) * np.dtype(datatype).itemsize
        cval_len = start_pos + requested_byte_size
        if byte_size < cval_len:
            _raise_error("The size of the shared memory region is insufficient to provide numpy array with requested size")
        if cval_len == 0:
            return np.empty(shape, dtype=datatype)
        val_buf = ctypes.cast(ctypes.c_void_p(addr), ctypes.POINTER(ctypes.c_byte * cval_len))[0]
        val = np.frombuffer(val_buf, dtype=datatype, offset=start_pos)
        return np.reshape(val, shape)
    count = max(int(np.prod(shape)), 1)
    if _has_device() and byte_size - start_pos >= 4096:
        from .. import _object_array_from, deserialize_from_device
        from ..._device import current_device, device_buffer, stream_ptr, to_host

        dev = current_device()
        nbytes = byte_size - start_pos
        d = device_buffer(nbytes, dev)
        _raise_if_error(_lib.lib.TcbShmRegionCopyDevice(shm_handle, start_pos, d.data_ptr(), nbytes, dev, 0,
                                                        stream_ptr(dev)))
        offsets, payload = deserialize_from_device(d[:nbytes], nbytes, count, 0, dev)
        h_off = to_host(offsets.data_ptr(), offsets.numel() * 8, dev).view(np.int64)
        h_pay = to_host(payload.data_ptr(), payload.numel(), dev)
        return np.reshape(_object_array_from(h_off, h_pay), shape)
    val_buf = ctypes.cast(ctypes.c_void_p(addr), ctypes.POINTER(ctypes.c_byte * byte_size))[0]
    str_offset = start_pos
    strs = []
    for _ in range(count):
        l = struct.unpack_from("<I", val_buf, str_offset)[0]
        str_offset += 4
        sb = struct.unpack_from("<{}s".format(l), val_buf, str_offset)[0]
        str_offset += l
        strs.append(sb)
    val = np.empty(len(strs), dtype=object)
    val[:] = strs
    return np.reshape(val, shape)


def set_shared_memory_region_from_device(shm_handle, tensors, offset=0, device=None):
    """Copy CUDA tensors (contiguous) back to back into the region from ``offset`` by DMA; returns the
    number of bytes written. No counterpart in the reference (it goes through a numpy array)."""
    from ..._device import stream_ptr

    offset_current = int(offset)
    for t in tensors:
        if not t.is_contiguous():
            _raise_error("each tensor must be contiguous")
        dev = t.device.index if device is None else device
        nbytes = int(t.numel() * t.element_size())
        _raise_if_error(_lib.lib.TcbShmRegionCopyDevice(shm_handle, offset_current, t.data_ptr(), nbytes, dev, 1,
                                                        stream_ptr(dev)))
        offset_current += nbytes
    return offset_current - int(offset)


def get_contents_to_device(shm_handle, tensor, offset=0, device=None):
    """Fill the (contiguous) CUDA ``tensor`` with the region's bytes from ``offset`` by DMA."""
    from ..._device import stream_ptr

    if not tensor.is_contiguous():
        _raise_error("the tensor must be contiguous")
    dev = tensor.device.index if device is None else device
    _raise_if_error(_lib.lib.TcbShmRegionCopyDevice(shm_handle, int(offset), tensor.data_ptr(),
                                                    int(tensor.numel() * tensor.element_size()), dev, 0,
                                                    stream_ptr(dev)))
    return tensor


def mapped_shared_memory_regions():
    """Keys of all regions created and not yet destroyed."""
    return mapped_shm_regions


def destroy_shared_memory_region(shm_handle):
    """munmap + shm_unlink (reference :274-311)."""
    _, key, _, _, _ = _info(shm_handle)
    mapped_shm_regions.remove(key)
    _raise_if_error(_lib.lib.TcbShmRegionDestroy(shm_handle), libcshm_codes=True)
    return
