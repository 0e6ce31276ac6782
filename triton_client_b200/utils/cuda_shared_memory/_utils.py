"""Region object and exception of the CUDA shared-memory module (reference:
tritonclient/utils/cuda_shared_memory/_utils.py:49-100)."""
import ctypes

from ... import _lib


class CudaSharedMemoryException(Exception):
    """Exception indicating non-Success status."""

    def __init__(self, msg):
        self._msg = msg

    def __str__(self):
        return super().__str__() if self._msg is None else self._msg


class _IpcHandle:
    """Stands in for cudart.cudaIpcMemHandle_t: `.reserved` is the 64 raw bytes
    (read at cuda_shared_memory/__init__.py:170)."""

    def __init__(self, raw):
        self.reserved = bytes(raw)


class CudaSharedMemoryRegion:
    """Owns one cudaMalloc allocation + its IPC handle through libtcb200 (TcbRegionCreate /
    TcbRegionDestroy). Same attributes as the reference's class: _triton_shm_name,
    _cuda_shm_handle, _base_addr, _byte_size, _device_id."""

    def __init__(self, triton_shm_name, byte_size, device_id):
        handle = ctypes.c_void_p()
        _lib.check(_lib.lib.TcbRegionCreate(triton_shm_name.encode("utf-8"), int(byte_size),
                                            int(device_id), ctypes.byref(handle)))
        base = ctypes.c_uint64()
        size = ctypes.c_uint64()
        dev = ctypes.c_int()
        raw = (ctypes.c_ubyte * 64)()
        _lib.check(_lib.lib.TcbRegionInfo(handle, ctypes.byref(base), ctypes.byref(size),
                                          ctypes.byref(dev), raw))
        self._c_handle = handle
        self._triton_shm_name = triton_shm_name
        self._cuda_shm_handle = _IpcHandle(raw)
        self._base_addr = int(base.value)
        self._byte_size = int(size.value)
        self._device_id = int(dev.value)

    def __del__(self):
        handle = getattr(self, "_c_handle", None)
        if handle:
            self._c_handle = None
            try:
                _lib.lib.TcbRegionDestroy(handle)
            except Exception:
                pass
