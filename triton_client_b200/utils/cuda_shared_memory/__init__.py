"""Drop-in for ``tritonclient.utils.cuda_shared_memory`` on a B200.

Same public functions, argument order and fixed error messages as
/root/reference/src/python/library/tritonclient/utils/cuda_shared_memory/__init__.py
(:107 create, :152 get_raw_handle, :173 set, :242 get_contents_as_numpy, :328 set_from_dlpack,
:391 as_shared_memory_tensor, :402 allocated_shared_memory_regions, :414 destroy); the CUDA work
goes through libtcb200.so's C ABI (``include/tcb200.h``) instead of cuda-python, and BYTES
contents are decoded on the device instead of by a Python ``unpack_from`` loop.

Additions that the reference API cannot express (all optional keyword arguments, so existing
call sites keep working): ``offset=`` on set/get, and device tensors accepted by
``set_shared_memory_region``.
"""
import base64
import ctypes
import struct

import numpy as np

from ... import _lib
from ..._device import stream_ptr, to_host
from .. import _dlpack
from .._shared_memory_tensor import SharedMemoryTensor
from . import _host_pool
from ._utils import CudaSharedMemoryException, CudaSharedMemoryRegion

allocated_shm_regions = []


def _get_or_create_global_cuda_stream(device_id):
    """One process-global stream per device (reference :66-70), owned by the library."""
    s = ctypes.c_void_p()
    _lib.check(_lib.lib.TcbGlobalStream(int(device_id), ctypes.byref(s)))
    return s.value


def create_shared_memory_region(triton_shm_name, byte_size, device_id):
    """cudaMalloc + cudaIpcGetMemHandle on ``device_id``; the returned object owns the memory."""
    try:
        region = CudaSharedMemoryRegion(triton_shm_name, byte_size, device_id)
    except Exception as ex:
        raise CudaSharedMemoryException("unable to create cuda shared memory handle") from ex
    allocated_shm_regions.append(region)
    return region


def get_raw_handle(cuda_shm_handle):
    """base64 of the 64-byte cudaIpcMemHandle_t, as consumed by register_cuda_shared_memory."""
    return base64.b64encode(cuda_shm_handle._cuda_shm_handle.reserved)


def set_shared_memory_region(cuda_shm_handle, input_values, offset=0):
    """Copy the arrays back to back into the region (from ``offset``, default 0 like the
    reference), one stream synchronisation at the end. ``np.object_`` arrays are expected to be
    the result of ``serialize_byte_tensor``."""
    if not isinstance(input_values, (list, tuple)):
        raise CudaSharedMemoryException("input_values must be specified as a numpy array")
    for input_value in input_values:
        if not isinstance(input_value, (np.ndarray,)):
            raise CudaSharedMemoryException(
                "input_values must be specified as a list/tuple of numpy arrays")
    try:
        dev = cuda_shm_handle._device_id
        stream = _get_or_create_global_cuda_stream(dev)
        offset_current = int(offset)
        keep = []  # host buffers must outlive the asynchronous copies
        for input_value in input_values:
            # the reference's ascontiguousarray(...).flatten() copies the array a second time; a flat
            # view of the contiguous array holds the same bytes
            input_value = np.ascontiguousarray(input_value).reshape(-1)
            if input_value.dtype == np.object_:
                data = input_value.item()
                src = np.frombuffer(data, dtype=np.uint8)
            else:
                src = input_value.view(np.uint8) if input_value.size else np.empty(0, np.uint8)
            keep.append(src)
            _lib.check(_lib.lib.TcbRegionSet(cuda_shm_handle._c_handle, offset_current,
                                             src.ctypes.data if src.size else None, src.size, stream))
            offset_current += src.size
        _lib.check(_lib.lib.TcbStreamSync(dev, stream))
    except Exception as ex:
        if not isinstance(ex, CudaSharedMemoryException):
            raise CudaSharedMemoryException("unable to set values in cuda shared memory") from ex
        raise ex
    return


def get_contents_as_numpy(cuda_shm_handle, datatype, shape, offset=0):
    """New numpy array with the region's contents interpreted as ``datatype``/``shape``.
    Fixed-size dtypes copy only the bytes asked for; BYTES (``np.object_`` / ``np.bytes_``) is
    decoded on the device (chunked offset chase + gather) and only the payload comes back."""
    dev = cuda_shm_handle._device_id
    shape = [int(s) for s in shape]
    start_pos = int(offset)
    if (datatype != np.object_) and (datatype != np.bytes_):
        requested_byte_size = int(np.prod(shape)) * np.dtype(datatype).itemsize
        cval_len = start_pos + requested_byte_size
        if cuda_shm_handle._byte_size < cval_len:
            raise CudaSharedMemoryException(
                "The size of the shared memory region is insufficient to provide numpy array with requested size")
        if requested_byte_size == 0:
            return np.empty(shape, dtype=datatype)
        # large results land in a recycled page-locked buffer (see _host_pool): no first-touch page
        # faults, no staging - the DMA engine writes the array the caller gets
        result, host_ptr = _host_pool.leased_array(requested_byte_size, datatype, shape)
        if result is None:
            result = np.empty(shape, dtype=datatype)
            host_ptr = result.ctypes.data
        try:
            stream = _get_or_create_global_cuda_stream(dev)
            _lib.check(_lib.lib.TcbRegionGet(cuda_shm_handle._c_handle, start_pos, host_ptr,
                                             requested_byte_size, stream))
        except Exception as ex:
            raise CudaSharedMemoryException("failed to read cuda shared memory results") from ex
        return result
    # BYTES: the reference reads prod(shape) elements but never fewer than one (:312)
    from .. import _object_array_from, deserialize_from_device

    count = max(int(np.prod(shape)), 1)
    try:
        offsets, payload = deserialize_from_device(cuda_shm_handle, None, count, start_pos, dev)
        h_off = to_host(offsets.data_ptr(), offsets.numel() * 8, dev).view(np.int64)
        h_pay = to_host(payload.data_ptr(), payload.numel(), dev)
    except struct.error:
        raise
    except Exception as ex:
        raise CudaSharedMemoryException("failed to read cuda shared memory results") from ex
    val = _object_array_from(h_off, h_pay)
    return np.reshape(val, shape)


def _is_device_supported(device_type):
    return int(device_type) in (_dlpack.kDLCPU, _dlpack.kDLCUDA, _dlpack.kDLCUDAHost)


def set_shared_memory_region_from_dlpack(cuda_shm_handle, input_values, offset=0):
    """Copy DLPack-exporting tensors (CPU, CUDA or pinned host; C-contiguous) back to back into
    the region. torch parses the capsules (buffer plumbing); the copy itself is TcbRegionSet."""
    import torch

    offset_current = int(offset)
    dev = cuda_shm_handle._device_id
    for input_value in input_values:
        if hasattr(input_value, "__dlpack_device__"):
            dl_device = input_value.__dlpack_device__()
            if not _is_device_supported(dl_device[0]):
                raise CudaSharedMemoryException(
                    "DLPack device type {} is not supported".format(dl_device[0]))
        t = input_value if isinstance(input_value, torch.Tensor) else torch.from_dlpack(input_value)
        if not t.is_contiguous():
            raise CudaSharedMemoryException(
                "DLPack tensor is not contiguous. Only contiguous DLPack tensors that are stored in C-Order are supported.")
        byte_size = t.numel() * t.element_size()
        try:
            # order the copy after the producer: work queued on torch's stream of that device
            if t.is_cuda:
                torch.cuda.current_stream(t.device).synchronize()
            stream = _get_or_create_global_cuda_stream(dev)
            _lib.check(_lib.lib.TcbRegionSet(cuda_shm_handle._c_handle, offset_current,
                                             t.data_ptr() if byte_size else None, byte_size, stream))
            _lib.check(_lib.lib.TcbStreamSync(dev, stream))
        except Exception as ex:
            if not isinstance(ex, CudaSharedMemoryException):
                raise CudaSharedMemoryException("unable to set values in cuda shared memory") from ex
            raise ex
        offset_current += byte_size
    return


def as_shared_memory_tensor(cuda_shm_handle, datatype, shape):
    return SharedMemoryTensor(datatype, shape, cuda_shm_handle._base_addr, 0,
                              cuda_shm_handle._byte_size, cuda_shm_handle._device_id,
                              owner=cuda_shm_handle)


def allocated_shared_memory_regions():
    """All regions created and not yet destroyed."""
    return allocated_shm_regions


def destroy_shared_memory_region(cuda_shm_handle):
    """Remove from the live list and drop the reference; the region object frees the device
    memory when its last reference goes (as in the reference, :414-429)."""
    allocated_shm_regions.remove(cuda_shm_handle)
    del cuda_shm_handle
    return
