"""CPU tests of the boundary: the C-ABI library loads, exports every symbol include/tcb200.h
declares, and the host-side logic (coefficient set-up, value tables, element normalisation,
argument validation) agrees with the oracle. No GPU compute is called here."""
import ctypes
import os
import re

import numpy as np
import pytest

from oracle import bytes_codec as bc
from oracle import clib
from oracle import pil_bilinear as pb
from tests.golden import cases

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "tcb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(Tcb[A-Z]\w+)\s*\(", text)))


def test_library_loads_and_exports_every_declared_symbol():
    from triton_client_b200 import _lib

    declared = _declared_symbols()
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(_lib.lib, name), "libtcb200.so does not export %s" % name
    assert sorted(_lib.EXPORTS) == declared
    assert _lib.version().startswith("tcb200")


def test_missing_library_fails_loudly(tmp_path, monkeypatch):
    import importlib
    import sys

    from triton_client_b200 import _lib

    real = _lib.LIB_PATH
    hidden = real + ".hidden"
    os.rename(real, hidden)
    try:
        saved = {k: sys.modules.pop(k) for k in list(sys.modules) if k.startswith("triton_client_b200")}
        with pytest.raises(ImportError, match="no CPU fallback"):
            importlib.import_module("triton_client_b200")
    finally:
        os.rename(hidden, real)
        for k in [k for k in sys.modules if k.startswith("triton_client_b200")]:
            sys.modules.pop(k)
        sys.modules.update(saved)


@pytest.mark.parametrize("pair", [(640, 224), (480, 224), (1920, 224), (1080, 224), (3840, 224),
                                  (2160, 224), (150, 224), (100, 224), (224, 224), (7, 224), (1000, 224),
                                  (1, 5), (5, 1), (299, 224), (225, 224), (4000, 299), (53, 299)])
def test_host_coefficients_equal_oracle(pair):
    from triton_client_b200 import _lib

    i, o = pair
    b_ref, k_ref, ks_ref = clib.precompute_coeffs(i, o)
    b = np.zeros((o, 2), dtype=np.int32)
    k = np.zeros((o, ks_ref), dtype=np.int32)
    ks = _lib.lib.TcbResampleCoeffs(i, o, b.ctypes.data, k.ctypes.data, k.size)
    assert ks == ks_ref
    assert np.array_equal(b, b_ref) and np.array_equal(k, k_ref)
    # capacity check
    assert _lib.lib.TcbResampleCoeffs(i, o, b.ctypes.data, k.ctypes.data, k.size - 1) == _lib.TCB_ERR_OUT_OF_RANGE
    assert "needs" in _lib.last_error()


@pytest.mark.parametrize("dtype", ["FP32", "FP16", "UINT8", "INT8", "INT16", "INT32", "UINT16", "FP64", "BF16", "NOPE"])
@pytest.mark.parametrize("scaling", ["NONE", "INCEPTION", "VGG", "other"])
@pytest.mark.parametrize("c", [1, 3])
def test_value_table_equals_oracle(dtype, scaling, c):
    from triton_client_b200.image_client import scaling_table

    try:
        b = pb.scaling_lut(dtype, c, scaling)
    except Exception as ex:  # e.g. INT8 + VGG + c==1: numpy itself refuses 128 as int8
        with pytest.raises(type(ex)):
            scaling_table(dtype, c, scaling)
        return
    a = scaling_table(dtype, c, scaling)
    assert a.dtype == b.dtype and a.shape == (c, 256)
    assert a.tobytes() == b.tobytes()


def test_dtype_maps_follow_reference():
    from triton_client_b200 import utils as u

    for name in ["BOOL", "INT8", "INT16", "INT32", "INT64", "UINT8", "UINT16", "UINT32", "UINT64",
                 "FP16", "FP32", "FP64", "BYTES", "BF16", "nope"]:
        assert u.triton_to_np_dtype(name) == pb.triton_to_np_dtype(name)
    assert u.np_to_triton_dtype(np.float32) == "FP32"
    assert u.np_to_triton_dtype(bool) == "BOOL"
    assert u.np_to_triton_dtype(np.dtype("S7")) == "BYTES"
    assert u.np_to_triton_dtype(np.dtype(object)) == "BYTES"
    assert u.np_to_triton_dtype(np.dtype("complex64")) is None


def test_element_normalisation_equals_oracle():
    from triton_client_b200 import utils as u

    for name, arr in cases.object_cases().items():
        if arr.dtype == np.object_ and arr.size:
            assert u._elements_as_bytes(arr) == bc.element_bytes(arr), name
    arr = np.array([b"ab", b"c"], dtype=np.object_)
    assert u.serialized_byte_size(arr) == 3
    with pytest.raises(u.InferenceServerException, match="must be np.object_"):
        u.serialized_byte_size(np.zeros(3))
    with pytest.raises(u.InferenceServerException, match="invalid datatype"):
        u.serialize_byte_tensor(np.zeros(3, dtype=np.float32))
    e = u.serialize_byte_tensor(np.empty((0,), dtype=np.object_))
    assert e.shape == (0,) and e.dtype == np.object_


def test_argument_validation_without_gpu():
    """Calls that must be rejected before any CUDA work happens."""
    from triton_client_b200 import _lib

    L = _lib.lib
    plan = ctypes.c_void_p()
    img = (_lib.TcbImage * 1)(_lib.TcbImage(0, 4, 4, 12, 3))
    lut = np.zeros((3, 256), dtype=np.float32)
    args = lambda **kw: [kw.get("dev", 0), kw.get("imgs", img), kw.get("n", 1), kw.get("c", 3), kw.get("h", 8),
                         kw.get("w", 8), kw.get("layout", 2), lut.ctypes.data, kw.get("elem", 4), kw.get("path", 0),
                         ctypes.byref(plan)]
    assert L.TcbPreprocessPlanCreate(*args(c=2)) == _lib.TCB_ERR_INVALID_ARG
    assert L.TcbPreprocessPlanCreate(*args(layout=0)) == _lib.TCB_ERR_INVALID_ARG
    assert L.TcbPreprocessPlanCreate(*args(elem=3)) == _lib.TCB_ERR_INVALID_ARG
    assert L.TcbPreprocessPlanCreate(*args(h=0)) == _lib.TCB_ERR_INVALID_ARG
    assert L.TcbPreprocessPlanCreate(*args(path=4)) == _lib.TCB_ERR_INVALID_ARG
    bad = (_lib.TcbImage * 1)(_lib.TcbImage(0, 4, 4, 11, 3))
    assert L.TcbPreprocessPlanCreate(*args(imgs=bad)) == _lib.TCB_ERR_INVALID_ARG
    assert "pitch" in _lib.last_error()
    bad = (_lib.TcbImage * 1)(_lib.TcbImage(0, 4, 4, 8, 2))
    assert L.TcbPreprocessPlanCreate(*args(imgs=bad)) == _lib.TCB_ERR_INVALID_ARG
    assert L.TcbRegionInfo(None, None, None, None, None) == _lib.TCB_ERR_INVALID_ARG
    assert L.TcbBytesSerializeWorkspace(0) > 0 and L.TcbBytesSerializeWorkspace(10**6) >= 8 * 977
    assert L.TcbBytesDeserializeWorkspace(1 << 20) > (1 << 20) // 32  # ~1 KiB of header list per 16 KiB chunk
    with pytest.raises(_lib.TcbError):
        _lib.check(-3)
