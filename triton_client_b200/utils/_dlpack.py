"""Minimal DLPack producer in ctypes, so that a CUDA shared-memory region can be handed to
``torch.from_dlpack`` without copying (behaviour of the reference's
tritonclient/utils/_dlpack.py:57-261 and _shared_memory_tensor.py:34-87; struct layouts are those
of dlpack.h v0.8).

Design: one keep-alive registry instead of manual reference counting. ``make_capsule`` parks every
ctypes object the consumer may still read (the DLManagedTensor, its shape array) in ``_LIVE`` under
the struct's address; the DLPack deleter (called by the consumer) or the capsule destructor (if
nobody consumed it) removes the entry.
"""
import ctypes

kDLCPU, kDLCUDA, kDLCUDAHost = 1, 2, 3
kDLInt, kDLUInt, kDLFloat, kDLBfloat, kDLBool = 0, 1, 2, 4, 6

_CAPSULE_NAME = b"dltensor"


class DLDevice(ctypes.Structure):
    _fields_ = [("device_type", ctypes.c_int), ("device_id", ctypes.c_int)]


class DLDataType(ctypes.Structure):
    _fields_ = [("code", ctypes.c_uint8), ("bits", ctypes.c_uint8), ("lanes", ctypes.c_uint16)]


class DLTensor(ctypes.Structure):
    _fields_ = [
        ("data", ctypes.c_void_p),
        ("device", DLDevice),
        ("ndim", ctypes.c_int),
        ("dtype", DLDataType),
        ("shape", ctypes.POINTER(ctypes.c_int64)),
        ("strides", ctypes.POINTER(ctypes.c_int64)),
        ("byte_offset", ctypes.c_uint64),
    ]


_DELETER = ctypes.CFUNCTYPE(None, ctypes.c_void_p)


class DLManagedTensor(ctypes.Structure):
    _fields_ = [("dl_tensor", DLTensor), ("manager_ctx", ctypes.c_void_p), ("deleter", _DELETER)]


_TRITON_TO_DL = {
    "BOOL": (kDLBool, 8), "INT8": (kDLInt, 8), "INT16": (kDLInt, 16), "INT32": (kDLInt, 32),
    "INT64": (kDLInt, 64), "UINT8": (kDLUInt, 8), "UINT16": (kDLUInt, 16), "UINT32": (kDLUInt, 32),
    "UINT64": (kDLUInt, 64), "FP16": (kDLFloat, 16), "FP32": (kDLFloat, 32), "FP64": (kDLFloat, 64),
    "BF16": (kDLBfloat, 16),
}


def triton_to_dlpack_dtype(dtype):
    if dtype == "BYTES":
        raise Exception("DLPack currently doesn't suppose BYTES type")
    if dtype not in _TRITON_TO_DL:
        raise Exception("Can not covert unknown data type '{}' to DLPack data type".format(dtype))
    code, bits = _TRITON_TO_DL[dtype]
    return DLDataType(code, bits, 1)


_LIVE = {}

_api = ctypes.pythonapi
_api.PyCapsule_New.restype = ctypes.py_object
_api.PyCapsule_New.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_void_p]
_api.PyCapsule_IsValid.restype = ctypes.c_int
_api.PyCapsule_IsValid.argtypes = [ctypes.py_object, ctypes.c_char_p]
_api.PyCapsule_GetPointer.restype = ctypes.c_void_p
_api.PyCapsule_GetPointer.argtypes = [ctypes.py_object, ctypes.c_char_p]


@_DELETER
def _managed_deleter(address):
    _LIVE.pop(address, None)


@ctypes.CFUNCTYPE(None, ctypes.c_void_p)
def _capsule_destructor(capsule_ptr):
    capsule = ctypes.cast(capsule_ptr, ctypes.py_object)
    # still named "dltensor" means no consumer took ownership: release our keep-alive entry
    if _api.PyCapsule_IsValid(capsule, _CAPSULE_NAME):
        _LIVE.pop(_api.PyCapsule_GetPointer(capsule, _CAPSULE_NAME), None)


def make_capsule(data_ptr, byte_offset, device, dtype, shape, owner=None):
    """PyCapsule("dltensor") describing a compact row-major view. `device` = (type, id)."""
    shape = [int(s) for s in shape]
    shape_arr = (ctypes.c_int64 * max(len(shape), 1))(*shape)
    managed = DLManagedTensor()
    t = managed.dl_tensor
    t.data = int(data_ptr)
    t.device = DLDevice(int(device[0]), int(device[1]))
    t.ndim = len(shape)
    t.dtype = triton_to_dlpack_dtype(dtype)
    t.shape = ctypes.cast(shape_arr, ctypes.POINTER(ctypes.c_int64))
    t.strides = ctypes.POINTER(ctypes.c_int64)()  # NULL: compact, row-major
    t.byte_offset = int(byte_offset)
    managed.manager_ctx = None
    managed.deleter = _managed_deleter
    address = ctypes.addressof(managed)
    _LIVE[address] = (managed, shape_arr, owner)
    return _api.PyCapsule_New(address, _CAPSULE_NAME, _capsule_destructor)
