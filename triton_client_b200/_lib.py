"""ctypes binding of libtcb200.so (the C ABI declared in include/tcb200.h).

Same loading style as the reference's libcshm binding
(/root/reference/src/python/library/tritonclient/utils/shared_memory/__init__.py:48-72):
``cdll.LoadLibrary`` + explicit ``restype``/``argtypes`` per export, integer return codes mapped
to exceptions. There is NO fallback: if the CUDA library is missing this module raises at import
time, and every product entry point goes through it.
"""
import ctypes
import os
from ctypes import POINTER, c_char_p, c_int, c_int32, c_int64, c_uint32, c_uint64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libtcb200.so")

TCB_OK = 0
TCB_ERR_INVALID_ARG = -1
TCB_ERR_CUDA = -2
TCB_ERR_OUT_OF_RANGE = -3
TCB_ERR_TRUNCATED = -4
TCB_ERR_UNSUPPORTED = -5
TCB_ERR_NOMEM = -6
TCB_ERR_INTERNAL = -7

LAYOUT_NHWC = 1
LAYOUT_NCHW = 2

# every symbol include/tcb200.h declares (tests check that the library exports all of them)
EXPORTS = [
    "TcbLastError", "TcbVersion", "TcbKernelLaunchCount",
    "TcbRegionCreate", "TcbRegionInfo", "TcbRegionSet", "TcbRegionGet", "TcbRegionDestroy",
    "TcbGlobalStream", "TcbStreamSync", "TcbHostAlloc", "TcbHostFree", "TcbDeviceAlloc",
    "TcbDeviceFree", "TcbMemcpy",
    "TcbPreprocessPlanCreate", "TcbPreprocessPlanRun", "TcbPreprocessPlanInfo",
    "TcbPreprocessPlanDestroy", "TcbResampleCoeffs",
    "TcbBytesSerializeWorkspace", "TcbBytesDeserializeWorkspace", "TcbBytesSerialize",
    "TcbBytesSerializeFixedWidth", "TcbBytesDeserialize", "TcbBytesSerializeNoAlloc",
    "TcbBf16Serialize", "TcbBf16Deserialize", "TcbBytesDeserializeCount", "TcbBytesDeserializeFill",
    "TcbShmRegionCreate", "TcbShmRegionSet", "TcbShmRegionInfo", "TcbShmRegionCopyDevice", "TcbShmRegionDestroy",
    "TcbClassificationFields",
]


class TcbImage(ctypes.Structure):
    _fields_ = [
        ("src_dev", c_uint64),
        ("width", c_int32),
        ("height", c_int32),
        ("pitch", c_int32),
        ("src_channels", c_int32),
    ]


class TcbError(Exception):
    """A libtcb200 call returned a negative code."""

    def __init__(self, code, msg):
        super().__init__("libtcb200 error %d: %s" % (code, msg))
        self.code = code
        self.msg = msg


if not os.path.exists(LIB_PATH):
    raise ImportError(
        "triton_client_b200: %s is missing. This package has no CPU fallback; build the CUDA "
        "library first: `python -c 'import __graft_entry__ as g; g.build()'` or "
        "`make -C triton_client_b200/csrc`." % LIB_PATH
    )

lib = ctypes.CDLL(LIB_PATH)


def _sig(name, restype, argtypes):
    fn = getattr(lib, name)
    fn.restype = restype
    fn.argtypes = argtypes
    return fn


_sig("TcbLastError", c_char_p, [])
_sig("TcbVersion", c_char_p, [])
_sig("TcbKernelLaunchCount", c_uint64, [])
_sig("TcbRegionCreate", c_int, [c_char_p, c_uint64, c_int, POINTER(c_void_p)])
_sig("TcbRegionInfo", c_int, [c_void_p, POINTER(c_uint64), POINTER(c_uint64), POINTER(c_int), c_void_p])
_sig("TcbRegionSet", c_int, [c_void_p, c_uint64, c_void_p, c_uint64, c_void_p])
_sig("TcbRegionGet", c_int, [c_void_p, c_uint64, c_void_p, c_uint64, c_void_p])
_sig("TcbRegionDestroy", c_int, [c_void_p])
_sig("TcbGlobalStream", c_int, [c_int, POINTER(c_void_p)])
_sig("TcbStreamSync", c_int, [c_int, c_void_p])
_sig("TcbHostAlloc", c_int, [c_uint64, POINTER(c_void_p)])
_sig("TcbHostFree", c_int, [c_void_p])
_sig("TcbDeviceAlloc", c_int, [c_int, c_uint64, POINTER(c_void_p)])
_sig("TcbDeviceFree", c_int, [c_int, c_void_p])
_sig("TcbMemcpy", c_int, [c_int, c_void_p, c_void_p, c_uint64, c_void_p, c_int])
_sig("TcbPreprocessPlanCreate", c_int,
     [c_int, POINTER(TcbImage), c_int, c_int, c_int, c_int, c_int, c_void_p, c_int, c_int,
      POINTER(c_void_p)])
_sig("TcbPreprocessPlanRun", c_int, [c_void_p, POINTER(TcbImage), c_void_p, c_uint64, c_void_p])
_sig("TcbPreprocessPlanInfo", c_int, [c_void_p, POINTER(c_uint64), POINTER(c_uint64), POINTER(c_int)])
_sig("TcbPreprocessPlanDestroy", c_int, [c_void_p])
_sig("TcbResampleCoeffs", c_int, [c_int, c_int, c_void_p, c_void_p, c_int])
_sig("TcbBytesSerializeWorkspace", c_uint64, [c_uint64])
_sig("TcbBytesDeserializeWorkspace", c_uint64, [c_uint64])
_sig("TcbBytesSerialize", c_int,
     [c_int, c_void_p, c_void_p, c_int, c_uint64, c_void_p, c_uint64, c_void_p, c_uint64,
      c_void_p, POINTER(c_uint64), c_void_p])
_sig("TcbBytesSerializeNoAlloc", c_int,
     [c_int, c_void_p, c_void_p, c_int, c_uint64, c_void_p, c_uint64, POINTER(c_uint64), c_void_p])
_sig("TcbBytesSerializeFixedWidth", c_int,
     [c_int, c_void_p, c_uint32, c_uint64, c_void_p, c_uint64, c_void_p, c_uint64, c_void_p,
      POINTER(c_uint64), c_void_p])
_sig("TcbBytesDeserialize", c_int,
     [c_int, c_void_p, c_uint64, c_int64, c_void_p, c_uint64, c_void_p, c_uint64, c_void_p,
      c_uint64, POINTER(c_uint64), POINTER(c_uint64), c_void_p])
_sig("TcbShmRegionCreate", c_int, [c_char_p, c_char_p, c_uint64, c_int, POINTER(c_void_p)])
_sig("TcbShmRegionSet", c_int, [c_void_p, c_uint64, c_uint64, c_void_p])
_sig("TcbShmRegionInfo", c_int, [c_void_p, POINTER(c_void_p), POINTER(c_char_p), POINTER(c_int), POINTER(c_uint64),
                                 POINTER(c_uint64), POINTER(c_int)])
_sig("TcbShmRegionCopyDevice", c_int, [c_void_p, c_uint64, c_void_p, c_uint64, c_int, c_int, c_void_p])
_sig("TcbShmRegionDestroy", c_int, [c_void_p])
_sig("TcbBytesDeserializeCount", c_int,
     [c_int, c_void_p, c_uint64, c_int64, c_void_p, c_uint64, POINTER(c_uint64), POINTER(c_uint64), c_void_p])
_sig("TcbBytesDeserializeFill", c_int,
     [c_int, c_void_p, c_uint64, c_int64, c_void_p, c_uint64, c_void_p, c_uint64, c_void_p,
      c_uint64, POINTER(c_uint64), POINTER(c_uint64), c_void_p])
_sig("TcbBf16Serialize", c_int, [c_int, c_void_p, c_uint64, c_void_p, c_uint64, c_void_p])
_sig("TcbBf16Deserialize", c_int, [c_int, c_void_p, c_uint64, c_void_p, c_uint64, c_void_p])
_sig("TcbClassificationFields", c_int, [c_int, c_void_p, c_void_p, c_int64, c_void_p, c_void_p])


def last_error():
    return lib.TcbLastError().decode("utf-8", "replace")


def check(rc):
    """Raise TcbError for a negative return code."""
    if rc < 0:
        raise TcbError(rc, last_error())
    return rc


def version():
    return lib.TcbVersion().decode()


def kernel_launch_count():
    return int(lib.TcbKernelLaunchCount())
