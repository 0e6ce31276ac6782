"""GPU parity tests for the BF16 wire codec through the C ABI: bit-exact
against goldens generated from the unmodified reference (utils/__init__.py:279-348) and against
the oracle at sizes the reference's per-element loop cannot reach."""
import ctypes
import hashlib
import json
import os
import struct

import numpy as np
import pytest

from oracle import bytes_codec as bc
from tests.golden import cases

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")
BF = json.load(open(os.path.join(GOLD, "bf16_golden.json")))


def sha(b):
    return hashlib.sha256(b).hexdigest()


@pytest.mark.parametrize("rec", BF["cases"], ids=[c["name"] for c in BF["cases"]])
def test_bf16_dropin_matches_reference_golden(rec):
    from triton_client_b200 import utils as u

    arr = cases.bf16_cases()[rec["name"]]
    r = u.serialize_bf16_tensor(arr)
    assert list(r.shape) == rec["result_shape"] and str(r.dtype) == rec["result_dtype"]
    ser = r.item()
    assert isinstance(ser, bytes) and sha(ser) == rec["serialized_sha256"]
    back = u.deserialize_bf16_tensor(ser)
    assert list(back.shape) == rec["back_shape"] and str(back.dtype) == rec["back_dtype"]
    assert sha(np.ascontiguousarray(back).tobytes()) == rec["back_sha256"]


def test_bf16_dropin_edges():
    from triton_client_b200 import utils as u

    e = u.serialize_bf16_tensor(np.zeros((0, 3), dtype=np.float32))
    assert e.shape == (0,) and e.dtype == np.object_
    with pytest.raises(u.InferenceServerException, match="cannot serialize bf16 tensor: invalid datatype"):
        u.serialize_bf16_tensor(np.zeros(4, dtype=np.float64))
    assert u.deserialize_bf16_tensor(b"").shape == (0,)
    with pytest.raises(struct.error) as ei:
        u.deserialize_bf16_tensor(b"abcde")
    with pytest.raises(struct.error) as eo:
        bc.bf16_deserialize(b"abcde")
    assert str(ei.value) == str(eo.value)
    assert u.deserialize_bf16_tensor(bytearray(b"\x80\x3f")).tolist() == [[1.0]]
    assert u.deserialize_bf16_tensor(memoryview(b"\x80\x3f\x00\xc0")).tolist() == [[1.0], [-2.0]]


@pytest.mark.parametrize("n,src_off,dst_off", [(1, 0, 0), (7, 0, 0), (8, 0, 0), (9, 1, 0), (4099, 3, 1), (100003, 2, 5),
                                               (1 << 20, 0, 0), (1 << 20, 1, 3), ((1 << 22) + 5, 0, 4)])
def test_bf16_device_path_any_alignment(n, src_off, dst_off):
    """Device-resident path at every head/tail/alignment combination of the vector loops; raw
    random bit patterns so that NaN payloads, denormals and infinities all occur."""
    import torch

    from triton_client_b200 import utils as u

    rng = np.random.default_rng(n + 17 * src_off + dst_off)
    bits = rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)
    want = bc.bf16_serialize(bits.view(np.float32))
    src_buf = torch.zeros(n + 8, dtype=torch.float32, device="cuda")
    src = src_buf[src_off:src_off + n]
    src.copy_(torch.from_numpy(bits.view(np.int32)).cuda().view(torch.float32))
    out_buf = torch.full((2 * n + 64,), 0xA5, dtype=torch.uint8, device="cuda")
    out, nb = u.serialize_bf16_to_device(src, out_buf, 2 * dst_off)
    assert nb == 2 * n
    host = out_buf.cpu().numpy()
    assert host[2 * dst_off:2 * dst_off + 2 * n].tobytes() == want
    assert (host[:2 * dst_off] == 0xA5).all() and (host[2 * dst_off + 2 * n:] == 0xA5).all()  # nothing outside
    # and back: expand from the (possibly odd-aligned) wire position
    back_buf = torch.full((n + 8,), float("nan"), dtype=torch.float32, device="cuda")
    back = back_buf[src_off:src_off + n]
    u.deserialize_bf16_from_device(out_buf, 2 * n, 2 * dst_off, out=back)
    got = back.cpu().numpy().view(np.uint32)
    assert got.tobytes() == bc.bf16_deserialize(want).tobytes()
    assert torch.isnan(back_buf[:src_off]).all() and torch.isnan(back_buf[src_off + n:]).all()


def test_bf16_into_region_and_errors():
    import torch

    from triton_client_b200 import _lib
    from triton_client_b200 import utils as u
    from triton_client_b200.utils import cuda_shared_memory as cudashm

    n = 50000
    vals = torch.randn(n, device="cuda")
    region = cudashm.create_shared_memory_region("bf16_region", 2 * n + 16, 0)
    try:
        _, nb = u.serialize_bf16_to_device(vals, region, 16)
        wire = cudashm.get_contents_as_numpy(region, np.uint8, [2 * n + 16])[16:]
        assert wire.tobytes() == bc.bf16_serialize(vals.cpu().numpy())
        back = u.deserialize_bf16_from_device(region, 2 * n, 16)
        assert back.cpu().numpy().tobytes() == bc.bf16_deserialize(wire.tobytes()).tobytes()
        with pytest.raises(_lib.TcbError) as ei:  # capacity is checked, nothing is written past the region
            u.serialize_bf16_to_device(vals, region, 18)
        assert ei.value.code == _lib.TCB_ERR_OUT_OF_RANGE
        with pytest.raises(struct.error):
            u.deserialize_bf16_from_device(region, 2 * n - 1, 16)
        with pytest.raises(u.InferenceServerException):
            u.serialize_bf16_to_device(vals.double())
    finally:
        cudashm.destroy_shared_memory_region(region)
    rc = _lib.lib.TcbBf16Serialize(0, ctypes.c_void_p(vals.data_ptr() + 2), 4, ctypes.c_void_p(vals.data_ptr()), 8, None)
    assert rc == _lib.TCB_ERR_INVALID_ARG and "aligned" in _lib.last_error()
