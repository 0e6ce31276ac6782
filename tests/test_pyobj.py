"""The C loops over Python objects (triton_client_b200/_pyobj, csrc/pyobj.c) against the reference's
own per-element rules (utils/__init__.py:226-243, 262-276 of the reference). CPU only."""
import numpy as np
import pytest

from oracle import bytes_codec as bc
from triton_client_b200 import utils as u

pyobj = u._pyobj


def test_helper_is_built():
    import os
    import sysconfig

    if not os.path.exists(os.path.join(sysconfig.get_paths()["include"], "Python.h")):
        pytest.skip("no Python headers on this machine: the Python loops are used")
    assert pyobj is not None, "triton_client_b200/_pyobj*.so missing: run __graft_entry__.build()"


needs_helper = pytest.mark.skipif(pyobj is None, reason="_pyobj not built")


def _gather(arr, threads=1):
    flat = np.ascontiguousarray(arr.ravel(order="C"))
    lens = np.empty(flat.size, dtype=np.int64)
    total, h = pyobj.prepare(flat.ctypes.data, flat.size, lens.ctypes.data)
    dst = np.full(total + 64, 0xA5, dtype=np.uint8)
    assert pyobj.copy(h, dst.ctypes.data, total, threads) == total
    assert (dst[total:] == 0xA5).all()
    return dst[:total].tobytes(), lens


class Odd(bytes):
    pass


class Boom:
    def __str__(self):
        raise RuntimeError("no str for you")


@needs_helper
def test_gather_follows_the_reference_element_rules():
    elems = [b"abc", "héllo", 42, 3.5, Odd(b"xyz"), np.bytes_(b"nb"), bytearray(b"ba"), b"", None, "", b"\x00\x01"]
    arr = np.empty(len(elems), dtype=np.object_)
    for i, e in enumerate(elems):
        arr[i] = e
    want = [e if type(e) == bytes else str(e).encode("utf-8") for e in elems]  # noqa: E721 (the reference's test)
    blob, lens = _gather(arr)
    assert blob == b"".join(want)
    assert lens.tolist() == [len(w) for w in want]
    # and through the oracle's serialiser: same wire bytes as the restated reference
    assert bc.element_bytes(arr) == want
    assert bc.serialize_elements(arr) == b"".join(len(w).to_bytes(4, "little") + w for w in want)


@needs_helper
def test_gather_multi_dimensional_c_order_and_threads():
    rng = np.random.default_rng(1)
    n = 60_000
    lens = rng.integers(0, 300, n)
    items = [rng.integers(0, 256, int(k), dtype=np.uint8).tobytes() for k in lens]
    arr = np.empty(n, dtype=np.object_)
    arr[:] = items
    arr2 = arr.reshape(300, 200).T  # not C-contiguous: logical C order is what counts
    want = b"".join(arr2.ravel(order="C").tolist())
    for threads in (1, 4, 16):
        blob, got_lens = _gather(arr2, threads)
        assert blob == want
        assert got_lens.tolist() == [len(x) for x in arr2.ravel(order="C").tolist()]


@needs_helper
def test_gather_propagates_python_errors_and_checks_capacity():
    arr = np.empty(2, dtype=np.object_)
    arr[0], arr[1] = b"ok", Boom()
    lens = np.empty(2, dtype=np.int64)
    with pytest.raises(RuntimeError, match="no str for you"):
        pyobj.prepare(arr.ctypes.data, 2, lens.ctypes.data)
    arr[1] = b"fine"
    total, h = pyobj.prepare(arr.ctypes.data, 2, lens.ctypes.data)
    small = np.empty(3, dtype=np.uint8)
    with pytest.raises(ValueError):
        pyobj.copy(h, small.ctypes.data, 3, 1)


def test_scatter_makes_the_elements_the_reference_makes():
    rng = np.random.default_rng(2)
    n = 50_000
    lens = rng.integers(0, 260, n)
    lens[:5] = 0
    payload = rng.integers(0, 256, int(lens.sum()), dtype=np.uint8)
    offsets = np.concatenate(([0], np.cumsum(lens))).astype(np.int64)
    got = u._object_array_from(offsets, payload)
    raw = payload.tobytes()
    assert got.dtype == np.object_ and got.shape == (n,)
    assert all(type(x) is bytes for x in got[:100])
    assert got.tolist() == [raw[offsets[i]:offsets[i + 1]] for i in range(n)]
    # round trip through the oracle's codec
    assert bc.deserialize_stream(bc.serialize_elements(got)) == got.tolist()


@needs_helper
def test_scatter_releases_what_the_slots_held():
    import sys

    marker = object()
    before = sys.getrefcount(marker)
    arr = np.empty(3, dtype=np.object_)
    arr[:] = [marker, marker, marker]
    assert sys.getrefcount(marker) == before + 3
    pyobj.scatter(np.frombuffer(b"abcdef", dtype=np.uint8).ctypes.data, np.array([0, 1, 3, 6], dtype=np.int64).ctypes.data,
                  3, arr.ctypes.data)
    assert sys.getrefcount(marker) == before
    assert arr.tolist() == [b"a", b"bc", b"def"]
