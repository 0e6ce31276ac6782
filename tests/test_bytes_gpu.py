"""GPU parity tests for the BYTES codec through the C ABI: bit-exact against the goldens
generated from the unmodified reference, the reference's known-answer vectors, and the oracle."""
import hashlib
import json
import os
import struct

import numpy as np
import pytest

from oracle import bytes_codec as bc
from oracle import clib
from tests.golden import cases

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")
BYT = json.load(open(os.path.join(GOLD, "bytes_golden.json")))


def sha(b):
    return hashlib.sha256(b).hexdigest()


@pytest.mark.parametrize("rec", BYT["objects"], ids=[o["name"] for o in BYT["objects"]])
def test_serialize_dropin_object_cases(rec):
    from triton_client_b200 import utils as u

    arr = cases.object_cases()[rec["name"]]
    r = u.serialize_byte_tensor(arr)
    assert list(r.shape) == rec["result_shape"] and str(r.dtype) == rec["result_dtype"]
    ser = r.item() if r.size > 0 else b""
    assert ser.hex() == rec["serialized_hex"]
    if ser:
        assert u.serialized_byte_size(r) == len(ser)
        back = u.deserialize_bytes_tensor(ser)
        assert back.dtype == np.object_ and back.ndim == 1
        assert list(back) == bc.deserialize_stream(ser)


def test_known_answer_cc_client_test():
    # /root/reference/src/c++/tests/cc_client_test.cc:1929-1953, 2140-2153
    from triton_client_b200 import utils as u

    wire = bytes([2, 0, 0, 0, 1, 127, 2, 0, 0, 0, 2, 127])
    assert u.serialize_byte_tensor(np.array([b"\x01\x7f", b"\x02\x7f"], dtype=np.object_)).item() == wire
    assert list(u.deserialize_bytes_tensor(wire)) == [b"\x01\x7f", b"\x02\x7f"]


@pytest.mark.parametrize("rec", BYT["gbytes"], ids=[str(g["n"]) for g in BYT["gbytes"]])
def test_gbytes_device_path_bit_exact(rec):
    import torch

    from triton_client_b200 import utils as u

    blob, lens = cases.gbytes(rec["n"])
    d_blob = torch.from_numpy(blob).cuda()
    for ldt in (np.int32, np.int64):
        d_lens = torch.from_numpy(lens.astype(ldt)).cuda()
        out, total = u.serialize_to_device(d_blob, d_lens)
        assert total == rec["serialized_bytes"]
        ser = out[:total].cpu().numpy()
        assert sha(ser.tobytes()) == rec["serialized_sha256"]
    offsets, payload = u.deserialize_from_device(out, total)
    assert offsets.numel() == rec["n"] + 1
    assert np.array_equal(np.diff(offsets.cpu().numpy()), lens)
    assert np.array_equal(payload.cpu().numpy(), blob)
    # limited decode (get_contents_as_numpy semantics): first m elements only
    m = rec["n"] // 3 + 1
    off2, pay2 = u.deserialize_from_device(out, total, max_elements=m)
    assert off2.numel() == m + 1
    assert np.array_equal(off2.cpu().numpy(), offsets.cpu().numpy()[: m + 1])
    assert np.array_equal(pay2.cpu().numpy(), blob[: int(off2[-1])])


def test_serialize_dropin_gbytes_1000_and_unaligned_destination():
    import torch

    from triton_client_b200 import utils as u

    blob, lens = cases.gbytes(1000)
    arr = cases.to_object_array(blob, lens)
    ser = u.serialize_byte_tensor(arr).item()
    assert sha(ser) == BYT["gbytes"][1]["serialized_sha256"]
    back = u.deserialize_bytes_tensor(ser)
    assert np.array_equal(back, arr)
    # arbitrary destination / source misalignment exercises the byte paths
    d_blob_pad = torch.zeros(blob.size + 16, dtype=torch.uint8, device="cuda")
    d_lens = torch.from_numpy(lens).cuda()
    for src_mis in (0, 1, 2, 3, 5):
        d_blob = d_blob_pad[src_mis: src_mis + blob.size]
        d_blob.copy_(torch.from_numpy(blob))
        for dst_mis in (0, 1, 4, 7, 13):
            out = torch.full((len(ser) + 64,), 0xAB, dtype=torch.uint8, device="cuda")
            _, total = u.serialize_to_device(d_blob, d_lens, out, dst_mis)
            h = out.cpu().numpy()
            assert total == len(ser)
            assert h[dst_mis: dst_mis + total].tobytes() == ser, (src_mis, dst_mis)
            assert (h[:dst_mis] == 0xAB).all() and (h[dst_mis + total:] == 0xAB).all()
            offs, pay = u.deserialize_from_device(out, total, stream_offset=dst_mis)
            assert np.array_equal(pay.cpu().numpy(), blob)


def _roundtrip(lens, payload_kind="random", seed=0):
    import torch

    from triton_client_b200 import utils as u

    lens = np.asarray(lens, dtype=np.int64)
    rng = np.random.default_rng(seed)
    total = int(lens.sum())
    if payload_kind == "random":
        blob = rng.integers(0, 256, total, dtype=np.uint8)
    elif payload_kind == "zeros":
        blob = np.zeros(total, dtype=np.uint8)
    elif payload_kind == "ascii":
        blob = rng.integers(0x20, 0x7F, total, dtype=np.uint8)
    elif payload_kind == "small_ints":  # payloads that look like plausible length prefixes
        blob = np.zeros(total, dtype=np.uint8)
        blob[::4] = rng.integers(0, 40, blob[::4].size, dtype=np.uint8)
    ref = clib.bytes_serialize(blob, lens)
    out, tot = u.serialize_to_device(torch.from_numpy(blob).cuda(), torch.from_numpy(lens).cuda())
    assert tot == ref.size
    assert np.array_equal(out[:tot].cpu().numpy(), ref)
    offsets, payload = u.deserialize_from_device(out, tot)
    assert np.array_equal(np.diff(offsets.cpu().numpy()), lens)
    assert np.array_equal(payload.cpu().numpy(), blob)


@pytest.mark.parametrize("kind", ["random", "zeros", "ascii", "small_ints"])
def test_roundtrip_payload_kinds(kind):
    rng = np.random.default_rng(3)
    _roundtrip(rng.integers(0, 257, 20000), kind, seed=4)


def test_roundtrip_edge_lengths():
    _roundtrip([0])
    _roundtrip([1])
    _roundtrip([0] * 5000)                       # prefixes only
    _roundtrip([3, 0, 0, 1, 0, 2] * 1000)        # many elements per 16-byte chunk
    _roundtrip([100000, 5, 70000, 0, 9000])      # elements far longer than a deserialise chunk
    _roundtrip([1 << 20])                        # a single 1 MiB element
    _roundtrip(np.full(3000, 600))               # longer than the look-behind window (sweep path)
    _roundtrip(np.random.default_rng(1).integers(0, 5000, 3000))
    _roundtrip([2] * 40000, "zeros")             # more headers per chunk than the per-chunk list holds


@pytest.mark.parametrize("kind", ["random", "zeros", "small_ints"])
def test_roundtrip_long_elements_wide_vote_and_sweep(kind):
    """Elements longer than the 512-byte entry window: the first vote finds no header, the wider
    second-chance votes (8 KiB, 128 KiB) or the warp-parallel sweep must repair every chunk. Zeros
    and small integers are payloads that look like length prefixes (every guess may be wrong)."""
    rng = np.random.default_rng(11)
    _roundtrip(np.full(20000, 1024), kind, seed=1)                 # 20 MB of 1 KiB elements: 8 KiB window
    _roundtrip(rng.integers(513, 8000, 6000), kind, seed=2)        # mixed, all beyond the small window
    _roundtrip(np.full(700, 40000), kind, seed=3)                  # 128 KiB window
    _roundtrip(rng.integers(20000, 120000, 300), kind, seed=4)
    _roundtrip([300000, 12, 700000, 0, 0, 1 << 21, 5, 150000], kind, seed=5)   # beyond every window: sweep
    _roundtrip(np.concatenate([rng.integers(0, 257, 30000), np.full(50, 100000), rng.integers(0, 257, 30000),
                               np.full(3000, 2000)]), kind, seed=6)  # short / very long / short / medium


def test_truncated_streams_raise_struct_error():
    from triton_client_b200 import utils as u

    good = bc.serialize_elements(np.array([b"abc", b"de", b"f" * 9000, b"", b"zz"], dtype=np.object_))
    for cut in (1, 2, 3, 5, 6, 8, 12, 14, 5000, len(good) - 5, len(good) - 3, len(good) - 1):
        with pytest.raises(struct.error):
            bc.deserialize_stream(good[:cut])      # the reference's behaviour (oracle restatement)
        with pytest.raises(struct.error):
            u.deserialize_bytes_tensor(good[:cut])
    for cut in (7, 13, len(good) - 6, len(good) - 10):  # cuts on an element boundary are valid streams
        assert list(u.deserialize_bytes_tensor(good[:cut])) == bc.deserialize_stream(good[:cut])
    assert len(u.deserialize_bytes_tensor(good)) == 5
    assert u.deserialize_bytes_tensor(b"").shape == (0,)
    # a header that claims more than the buffer holds
    with pytest.raises(struct.error):
        u.deserialize_bytes_tensor(struct.pack("<I", 100) + b"x" * 10)


def test_serialize_into_region_and_capacity_error():
    import torch

    from triton_client_b200 import _lib
    from triton_client_b200 import utils as u
    from triton_client_b200.utils import cuda_shared_memory as cudashm

    blob, lens = cases.gbytes(1000)
    arr = cases.to_object_array(blob, lens)
    ref = bc.serialize_elements(arr)
    region = cudashm.create_shared_memory_region("ser", len(ref) + 100, 0)
    try:
        n = u.serialize_byte_tensor_to(region, arr, out_offset=32)
        assert n == len(ref)
        got = cudashm.get_contents_as_numpy(region, np.uint8, [len(ref)], offset=32)
        assert got.tobytes() == ref
        back = cudashm.get_contents_as_numpy(region, np.object_, [10, 100], offset=32)
        assert back.shape == (10, 100) and np.array_equal(back.reshape(-1), arr)
        # S<k> arrays: trailing NULs stripped on the device
        s_arr = np.array([b"\x00\x01\x02\x03", b"ab\x00\x00", b"\x00\x00\x00\x00", b"abcd"] * 300, dtype=np.bytes_)
        n2 = u.serialize_byte_tensor_to(region, s_arr)
        got2 = cudashm.get_contents_as_numpy(region, np.uint8, [n2])
        assert got2.tobytes() == bc.serialize_elements(s_arr)
        small = torch.zeros(100, dtype=torch.uint8, device="cuda")
        with pytest.raises(_lib.TcbError) as ei:
            u.serialize_to_device(torch.from_numpy(blob).cuda(), torch.from_numpy(lens).cuda(), small)
        assert ei.value.code == _lib.TCB_ERR_OUT_OF_RANGE
    finally:
        cudashm.destroy_shared_memory_region(region)


def test_full_size_cfg4_roundtrip_16m_strings():
    """cfg4 at full size: 16M strings of 0..256 bytes (~2.1 GB stream) generated and serialised on
    the device, then decoded from a CUDA shared-memory region. Size-independent properties:
    decode(serialize(x)) == x for lengths and payload, offsets are non-decreasing and end at the
    payload size, a limited decode returns a prefix, and the stream length equals P + 4N."""
    import torch

    from triton_client_b200 import utils as u
    from triton_client_b200.utils import cuda_shared_memory as cudashm

    N = 16_000_000
    g = torch.Generator(device="cuda").manual_seed(43)
    lens = torch.randint(0, 257, (N,), dtype=torch.int32, device="cuda", generator=g)
    P = int(lens.sum(dtype=torch.int64).item())
    blob = torch.randint(0, 256, (P,), dtype=torch.uint8, device="cuda", generator=g)
    region = cudashm.create_shared_memory_region("cfg4", P + 4 * N, 0)
    try:
        _, total = u.serialize_to_device(blob, lens, out=region)
        assert total == P + 4 * N
        offsets, payload = u.deserialize_from_device(region, total)
        assert offsets.numel() == N + 1 and payload.numel() == P
        assert torch.equal(offsets[1:] - offsets[:-1], lens.to(torch.int64))
        assert int(offsets[0]) == 0 and int(offsets[-1]) == P
        assert torch.equal(payload, blob)
        m = 1_234_567
        off2, pay2 = u.deserialize_from_device(region, None, max_elements=m)
        assert off2.numel() == m + 1 and torch.equal(off2, offsets[: m + 1])
        assert torch.equal(pay2, blob[: int(off2[-1])])
        # a strided sample of elements through the reference-facing API
        h = cudashm.get_contents_as_numpy(region, np.object_, [1000])
        lh = lens[:1000].cpu().numpy()
        bh = blob[: int(lh.sum())].cpu().numpy().tobytes()
        pos = 0
        for i in range(1000):
            assert h[i] == bh[pos: pos + lh[i]]
            pos += int(lh[i])
    finally:
        cudashm.destroy_shared_memory_region(region)


@pytest.mark.parametrize("case", ["mixed", "short", "long", "empties", "one_chunk", "chunk_edge"])
def test_decode_limits_garbage_tails_and_chunk_edges(case):
    """Offsets and payload must equal the oracle's for unlimited decodes, for limited decodes at
    every interesting cut, and with garbage after the requested elements (a region larger than
    its contents, as get_contents_as_numpy sees it); truncation must raise struct.error."""
    import torch

    from triton_client_b200 import utils as u

    rng = np.random.default_rng(11)
    lens = {
        "mixed": rng.integers(0, 257, 30000),
        "short": rng.integers(0, 6, 60000),
        "long": rng.integers(3000, 40000, 60),
        "empties": np.zeros(9000, dtype=np.int64),
        "one_chunk": rng.integers(0, 100, 50),
        # elements that end exactly on / one byte around the 16 KiB chunk boundaries
        "chunk_edge": np.array([16380, 16380, 16379, 1, 16376, 16385, 0, 16371, 4, 32764, 5], dtype=np.int64),
    }[case].astype(np.int64)
    total = int(lens.sum())
    blob = rng.integers(0, 256, total, dtype=np.uint8)
    ref = clib.bytes_serialize(blob, lens)
    ref_off = np.concatenate([[0], np.cumsum(lens)])
    tail = rng.integers(0, 256, 50000, dtype=np.uint8)
    for garbage in (None, "random", "zeros"):
        h = ref if garbage is None else np.concatenate([ref, tail if garbage == "random" else np.zeros_like(tail)])
        out = torch.from_numpy(h).cuda()
        cuts = [len(lens)] if garbage else [-1, len(lens)]
        cuts += [1, len(lens) // 2, len(lens) - 1]
        for m in cuts:
            oa, pa = u.deserialize_from_device(out, h.size, max_elements=m)
            n = len(lens) if m < 0 else m
            assert np.array_equal(oa.cpu().numpy(), ref_off[: n + 1]), (case, garbage, m)
            assert np.array_equal(pa.cpu().numpy(), blob[: ref_off[n]]), (case, garbage, m)
    out = torch.from_numpy(ref).cuda()
    for kw in ({"nbytes": ref.size - 1}, {"nbytes": ref.size, "max_elements": len(lens) + 1}):
        with pytest.raises(struct.error):
            u.deserialize_from_device(out, kw["nbytes"], max_elements=kw.get("max_elements", -1))


def test_tile_larger_than_2gib_takes_the_64bit_copy():
    """One tile of 1024 elements whose output exceeds 2^31 bytes: the copy falls back from 32-bit
    tile-relative offsets to 64-bit ones. Checked on the device slice by slice (header words at
    their known positions, payload slices equal to the blob) and by the round trip."""
    import torch

    from triton_client_b200 import utils as u

    big = (1 << 30) + (1 << 28) + 3
    lens_h = np.array([big, 5, big - 1, 0, 7, 13], dtype=np.int64)
    lens = torch.from_numpy(lens_h).cuda()
    P = int(lens_h.sum())
    g = torch.Generator(device="cuda").manual_seed(5)
    blob = torch.randint(0, 256, (P,), dtype=torch.uint8, device="cuda", generator=g)
    out, total = u.serialize_to_device(blob, lens)
    assert total == P + 4 * len(lens_h)
    pos = 0
    src = 0
    for n in lens_h:
        n = int(n)
        hdr = out[pos: pos + 4].cpu().numpy().tobytes()
        assert struct.unpack("<I", hdr)[0] == n
        assert torch.equal(out[pos + 4: pos + 4 + n], blob[src: src + n])
        pos += 4 + n
        src += n
    offsets, payload = u.deserialize_from_device(out, total)
    assert np.array_equal(np.diff(offsets.cpu().numpy()), lens_h)
    assert torch.equal(payload, blob)


def test_single_launch_path_for_streams_of_one_chunk():
    """Streams of at most 16 KiB decoded by TcbBytesDeserialize take k_des_tiny (one CTA, one launch):
    every cut point of a small stream, element limits, garbage tails, unaligned stream pointers and
    capacity errors against the oracle's walk (utils/__init__.py:262-276 / cuda_shared_memory:305-323)."""
    import ctypes

    import torch

    from triton_client_b200 import _lib

    rng = np.random.default_rng(77)
    ws_bytes = int(_lib.lib.TcbBytesDeserializeWorkspace(16384))
    ws = torch.empty(max(ws_bytes, 256), dtype=torch.uint8, device="cuda")

    def decode(stream, max_elements=-1, mis=0, ocap=None, pcap=None):
        buf = torch.empty(stream.size + 64, dtype=torch.uint8, device="cuda")
        buf[mis:mis + stream.size] = torch.from_numpy(stream).cuda()
        ocap = stream.size // 4 + 1 if ocap is None else ocap
        pcap = stream.size if pcap is None else pcap
        offs = torch.full((ocap + 2,), -7, dtype=torch.int64, device="cuda")
        pay = torch.full((pcap + 8,), 0xA5, dtype=torch.uint8, device="cuda")
        n, pb = ctypes.c_uint64(0), ctypes.c_uint64(0)
        before = _lib.kernel_launch_count()
        rc = _lib.lib.TcbBytesDeserialize(0, buf.data_ptr() + mis, stream.size, max_elements, offs.data_ptr(), ocap,
                                          pay.data_ptr(), pcap, ws.data_ptr(), int(ws.numel()), ctypes.byref(n),
                                          ctypes.byref(pb), None)
        launches = _lib.kernel_launch_count() - before
        return rc, int(n.value), int(pb.value), offs.cpu().numpy(), pay.cpu().numpy(), launches

    for trial in range(12):
        n = int(rng.integers(1, 120))
        lens = rng.integers(0, 90, n)
        lens[rng.integers(0, n)] = 0
        blob = rng.integers(0, 256, int(lens.sum()), dtype=np.uint8)
        stream = clib.bytes_serialize(blob, lens.astype(np.int64))
        assert stream.size <= 16384
        want_off = np.concatenate(([0], np.cumsum(lens)))
        rc, got_n, got_pb, offs, pay, launches = decode(stream, mis=trial % 5)
        assert rc == 0 and launches == 1
        assert got_n == n and got_pb == blob.size
        assert np.array_equal(offs[:n + 1], want_off) and (offs[n + 1:] == -7).all()
        assert np.array_equal(pay[:blob.size], blob) and (pay[blob.size:] == 0xA5).all()
        # element limit (get_contents_as_numpy semantics): m elements, whatever follows is ignored
        m = int(rng.integers(1, n + 1))
        rc, got_n, got_pb, offs, pay, _ = decode(np.concatenate((stream, rng.integers(0, 256, 9, dtype=np.uint8))), m)
        assert rc == 0 and got_n == m and got_pb == want_off[m]
        assert np.array_equal(offs[:m + 1], want_off[:m + 1]) and np.array_equal(pay[:got_pb], blob[:got_pb])
        # more elements asked for than the buffer holds: the reference unpacks past the end
        assert decode(stream, n + 1)[0] == _lib.TCB_ERR_TRUNCATED
        # every kind of cut: inside a prefix, inside a payload, exactly between elements
        for cut in sorted(set(rng.integers(1, stream.size, 6).tolist())):
            part = stream[:cut]
            try:
                w_blob, w_off = bc.deserialize_blob(part)
                ok = True
            except bc.OracleError:
                ok = False
            rc, got_n, got_pb, offs, pay, _ = decode(part)
            if ok:
                assert rc == 0 and got_n == w_off.size - 1 and np.array_equal(offs[:got_n + 1], w_off)
                assert np.array_equal(pay[:got_pb], w_blob)
            else:
                assert rc == _lib.TCB_ERR_TRUNCATED
        # capacities
        if n > 1:
            assert decode(stream, ocap=n - 1)[0] == _lib.TCB_ERR_OUT_OF_RANGE
        if blob.size > 1:
            rc, _, _, _, pay, _ = decode(stream, pcap=blob.size - 1)
            assert rc == _lib.TCB_ERR_OUT_OF_RANGE and (pay == 0xA5).all()
    # exactly one chunk, all-zero lengths (4096 empty elements)
    z = np.zeros(16384, dtype=np.uint8)
    rc, got_n, got_pb, offs, _, launches = decode(z)
    assert rc == 0 and launches == 1 and got_n == 4096 and got_pb == 0 and not offs[:4097].any()
