"""A DLPack view of a shared-memory region (reference: tritonclient/utils/
_shared_memory_tensor.py:34-87). The view is invalid once the region is modified or destroyed."""
from . import _dlpack


class SharedMemoryTensor:
    def __init__(self, dtype, shape, shm_addr, offset, byte_size, device_id, owner=None):
        self._dtype = dtype
        self._shape = shape
        self._shm_addr = shm_addr
        self._offset = offset
        self._byte_size = byte_size
        self._device_id = device_id
        self._owner = owner
        if device_id != -1:
            self._dl_device = (_dlpack.kDLCUDA, device_id)
        else:
            self._dl_device = (_dlpack.kDLCPU, 0)

    def __dlpack__(self, stream=None):
        return _dlpack.make_capsule(self._shm_addr, self._offset, self._dl_device, self._dtype,
                                    list(self._shape), owner=self._owner)

    def __dlpack_device__(self):
        return self._dl_device
