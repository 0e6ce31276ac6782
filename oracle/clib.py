"""Oracle (test infrastructure): ctypes binding of oracle/liboracle.so (built by oracle/Makefile)."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build():
    subprocess.check_call(["make", "-C", _HERE, "-s"])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        L = ctypes.CDLL(path)
        u8p = ctypes.c_void_p
        L.orc_ksize.restype = ctypes.c_int
        L.orc_ksize.argtypes = [ctypes.c_int, ctypes.c_int]
        L.orc_precompute_coeffs.restype = ctypes.c_int
        L.orc_precompute_coeffs.argtypes = [ctypes.c_int, ctypes.c_int, u8p, u8p]
        L.orc_resize_bilinear_u8.restype = ctypes.c_int
        L.orc_resize_bilinear_u8.argtypes = [u8p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                             ctypes.c_size_t, ctypes.c_int, ctypes.c_int, u8p]
        L.orc_rgb_to_l.restype = None
        L.orc_rgb_to_l.argtypes = [u8p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_size_t, u8p]
        L.orc_bytes_serialize.restype = ctypes.c_int64
        L.orc_bytes_serialize.argtypes = [u8p, u8p, ctypes.c_int64, u8p]
        L.orc_bytes_deserialize.restype = ctypes.c_int64
        L.orc_bytes_deserialize.argtypes = [u8p, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, u8p, u8p]
        _LIB = L
    return _LIB


def precompute_coeffs(in_size, out_size):
    L = lib()
    ks = L.orc_ksize(in_size, out_size)
    bounds = np.zeros((out_size, 2), dtype=np.int32)
    kk = np.zeros((out_size, ks), dtype=np.int32)
    L.orc_precompute_coeffs(in_size, out_size, bounds.ctypes.data, kk.ctypes.data)
    return bounds, kk, ks


def resize_bilinear_u8(src, out_w, out_h):
    src = np.ascontiguousarray(src)
    if src.ndim == 2:
        src = src[:, :, None]
    h, w, c = src.shape
    dst = np.empty((out_h, out_w, c), dtype=np.uint8)
    rc = lib().orc_resize_bilinear_u8(src.ctypes.data, h, w, c, w * c, out_w, out_h, dst.ctypes.data)
    if rc:
        raise MemoryError("orc_resize_bilinear_u8 failed")
    return dst


def rgb_to_l(src):
    src = np.ascontiguousarray(src)
    h, w, cin = src.shape
    dst = np.empty((h, w), dtype=np.uint8)
    lib().orc_rgb_to_l(src.ctypes.data, h, w, cin, w * cin, dst.ctypes.data)
    return dst


def bytes_serialize(blob, lens):
    blob = np.ascontiguousarray(blob, dtype=np.uint8)
    lens = np.ascontiguousarray(lens, dtype=np.int64)
    out = np.empty(int(lens.sum()) + 4 * lens.shape[0], dtype=np.uint8)
    n = lib().orc_bytes_serialize(blob.ctypes.data, lens.ctypes.data, lens.shape[0], out.ctypes.data)
    assert n == out.shape[0]
    return out


def bytes_deserialize(stream, max_elems=-1, capacity=None):
    stream = np.ascontiguousarray(stream, dtype=np.uint8)
    if capacity is None:
        capacity = stream.shape[0] // 4 + 1
    offsets = np.zeros(capacity + 1, dtype=np.int64)
    blob = np.empty(max(stream.shape[0], 1), dtype=np.uint8)
    n = lib().orc_bytes_deserialize(stream.ctypes.data, stream.shape[0], max_elems, capacity,
                                    offsets.ctypes.data, blob.ctypes.data)
    if n < 0:
        raise ValueError({-1: "truncated length prefix", -2: "truncated payload",
                          -3: "capacity exceeded"}[int(n)])
    return blob[: int(offsets[n])].copy(), offsets[: n + 1].copy()
