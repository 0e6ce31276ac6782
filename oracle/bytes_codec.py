"""Oracle (test infrastructure): CPU restatement of the BYTES tensor wire format.

Wire format (``/root/reference/src/python/library/tritonclient/utils/__init__.py:193-276``):
every element, in logical C order, is a 4-byte little-endian unsigned length followed by that
many payload bytes; elements are concatenated with no padding.

``serialize_elements`` / ``deserialize_stream`` are per-element loops that follow the reference
line by line (small cases only); ``serialize_blob`` / ``deserialize_blob`` are the vectorised
"lengths -> exclusive scan -> scatter" and "offset chase -> gather" statements that the CUDA
kernels mirror, operating on (payload blob, lengths) pairs so that large cases finish in
seconds.
"""
import struct

import numpy as np


class OracleError(Exception):
    pass


def element_bytes(input_tensor):
    """The per-element normalisation of utils/__init__.py:229-239 as a list of ``bytes`` in
    logical C order: object arrays keep exact ``bytes`` and ``str()``-encode everything else;
    ``np.bytes_`` (``S<k>``) arrays yield ``.item()`` (numpy strips trailing NULs)."""
    if (input_tensor.dtype != np.object_) and (input_tensor.dtype.type != np.bytes_):
        raise OracleError("cannot serialize bytes tensor: invalid datatype")
    out = []
    for obj in np.nditer(input_tensor, flags=["refs_ok"], order="C"):
        if input_tensor.dtype == np.object_:
            if type(obj.item()) == bytes:
                s = obj.item()
            else:
                s = str(obj.item()).encode("utf-8")
        else:
            s = obj.item()
        out.append(s)
    return out


def serialize_elements(input_tensor):
    """utils/__init__.py:193-246. Returns the serialized ``bytes`` (the reference wraps it in a
    0-d object array; empty input returns an empty 1-D object array there, ``b""`` here)."""
    if input_tensor.size == 0:
        return b""
    parts = []
    for s in element_bytes(input_tensor):
        parts.append(struct.pack("<I", len(s)))
        parts.append(s)
    return b"".join(parts)


def serialized_byte_size(tensor_value):
    """utils/__init__.py:43-68."""
    if tensor_value.dtype != np.object_:
        raise OracleError("The tensor_value dtype must be np.object_")
    total = 0
    if tensor_value.size > 0:
        for obj in np.nditer(tensor_value, flags=["refs_ok"], order="C"):
            total += len(obj.item())
    return total


def deserialize_stream(encoded, max_elements=None):
    """utils/__init__.py:249-276 (``max_elements=None``: run to the end of the buffer) and the
    inlined copy in cuda_shared_memory/__init__.py:305-323 (``max_elements=n``: stop after n).
    Raises ``struct.error`` on a truncated header or payload exactly as the reference does."""
    strs = []
    offset = 0
    buf = encoded
    while offset < len(buf):
        if max_elements is not None and len(strs) >= max_elements:
            break
        l = struct.unpack_from("<I", buf, offset)[0]
        offset += 4
        sb = struct.unpack_from("<{}s".format(l), buf, offset)[0]
        offset += l
        strs.append(sb)
    return strs


def cudashm_element_count(shape):
    """Elements read by get_contents_as_numpy's BYTES loop (cuda_shared_memory/__init__.py:312
    ``while (ii % np.prod(shape) != 0) or (ii == 0)``): prod(shape) elements, but never fewer
    than one - ``shape == []`` reads one element and returns a 0-d array; ``prod(shape) == 0``
    also reads one element and the final ``np.reshape`` then raises ValueError (probed)."""
    prod = int(np.prod(shape))
    return max(prod, 1)


# ----------------------------------------------------------------------------------------------
# Vectorised statements (what the kernels do)


def serialize_blob(blob, lengths):
    """blob: uint8[P] payloads back to back; lengths: int[N]. Returns uint8[P + 4N].

    offsets = exclusive_scan(lengths + 4); out[offsets[i] : +4] = LE32(lengths[i]);
    out[offsets[i]+4 : +lengths[i]] = payload i."""
    blob = np.asarray(blob, dtype=np.uint8)
    lengths = np.asarray(lengths, dtype=np.int64)
    n = lengths.shape[0]
    src_off = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(lengths, out=src_off[1:])
    assert src_off[-1] == blob.shape[0]
    dst_off = src_off + 4 * np.arange(n + 1, dtype=np.int64)
    out = np.empty(int(dst_off[-1]), dtype=np.uint8)
    # payload byte j of the blob belongs to element e(j) and lands at j + 4*(e(j)+1)
    elem_of_byte = np.repeat(np.arange(n, dtype=np.int64), lengths)
    out[np.arange(blob.shape[0], dtype=np.int64) + 4 * (elem_of_byte + 1)] = blob
    pref = lengths.astype("<u4").view(np.uint8).reshape(n, 4)
    idx = dst_off[:-1, None] + np.arange(4, dtype=np.int64)[None, :]
    out[idx] = pref
    return out


def deserialize_blob(stream, max_elements=None):
    """stream: uint8[S]. Returns (payload blob uint8[P], offsets int64[N+1]) with element i =
    blob[offsets[i]:offsets[i+1]]. Raises OracleError on truncation (the reference raises
    struct.error)."""
    stream = np.asarray(stream, dtype=np.uint8)
    s = stream.shape[0]
    mv = memoryview(stream)
    pos = 0
    starts = []
    lens = []
    unpack = struct.Struct("<I").unpack_from
    while pos < s:
        if max_elements is not None and len(lens) >= max_elements:
            break
        if pos + 4 > s:
            raise OracleError("truncated length prefix at offset %d" % pos)
        l = unpack(mv, pos)[0]
        pos += 4
        if pos + l > s:
            raise OracleError("truncated payload at offset %d" % pos)
        starts.append(pos)
        lens.append(l)
        pos += l
    n = len(lens)
    lens = np.asarray(lens, dtype=np.int64)
    starts = np.asarray(starts, dtype=np.int64)
    offsets = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(lens, out=offsets[1:])
    total = int(offsets[-1])
    if total:
        elem = np.repeat(np.arange(n, dtype=np.int64), lens)
        within = np.arange(total, dtype=np.int64) - offsets[elem]
        blob = stream[starts[elem] + within]
    else:
        blob = np.empty(0, dtype=np.uint8)
    return blob, offsets


def gbytes(n, seed=42, max_len=256):
    """The G-BYTES recipe of SURVEY.md 8c: lengths ~ U[0, max_len], payload uniform bytes."""
    rng = np.random.default_rng(seed)
    lens = rng.integers(0, max_len + 1, n, dtype=np.int64)
    blob = rng.integers(0, 256, int(lens.sum()), dtype=np.uint8)
    return blob, lens


def to_object_array(blob, lens):
    off = np.zeros(len(lens) + 1, dtype=np.int64)
    np.cumsum(lens, out=off[1:])
    b = blob.tobytes()
    arr = np.empty(len(lens), dtype=np.object_)
    for i in range(len(lens)):
        arr[i] = b[off[i]:off[i + 1]]
    return arr


# ---- BF16 wire codec (utils/__init__.py:279-348) ---------------------------------------------


def bf16_serialize_elements(input_tensor):
    """utils/__init__.py:279-323, per element: ``struct.pack("<f", x)[2:4]`` in logical C order.
    Returns the wire ``bytes`` (the reference wraps them in a 0-d object array)."""
    if input_tensor.size == 0:
        return b""
    if input_tensor.dtype != np.float32:
        raise OracleError("cannot serialize bf16 tensor: invalid datatype")
    parts = []
    for obj in np.nditer(input_tensor, flags=["refs_ok"], order="C"):
        parts.append(struct.pack("<f", obj)[2:4])
    return b"".join(parts)


def bf16_serialize(input_tensor):
    """Vectorised statement of the same thing: the upper 16 bits of every float32 pattern,
    little endian (truncation, no rounding). One subtlety, found by comparing with the reference on
    raw bit patterns: ``struct.pack("<f", obj)`` goes through a Python float (a double), and that
    round trip turns a signalling NaN into a quiet one - mantissa bit 22 is set before the cut, so
    0x7F800001 becomes 0x7FC0 and not the infinity 0x7F80."""
    bits = np.ascontiguousarray(input_tensor, dtype=np.float32).reshape(-1).view(np.uint32)
    nan = (bits & np.uint32(0x7FFFFFFF)) > np.uint32(0x7F800000)
    bits = np.where(nan, bits | np.uint32(0x00400000), bits)
    return (bits >> 16).astype("<u2").tobytes()


def bf16_deserialize(buf):
    """utils/__init__.py:326-348: ``b"\\x00\\x00" + two bytes`` -> float32. The reference appends
    the 1-tuples ``struct.unpack`` returns, hence the (n, 1) result shape (``(0,)`` when empty); an
    odd trailing byte raises ``struct.error``."""
    buf = bytes(buf)
    if len(buf) == 0:
        return np.array([], dtype=np.float32)
    if len(buf) & 1:
        raise struct.error("unpack_from requires a buffer of at least %d bytes for unpacking 2 bytes at offset %d "
                           "(actual buffer size is %d)" % (len(buf) + 1, len(buf) - 1, len(buf)))
    bits = np.frombuffer(buf, dtype="<u2").astype(np.uint32) << 16
    nan = (bits & np.uint32(0x7FFFFFFF)) > np.uint32(0x7F800000)
    bits = np.where(nan, bits | np.uint32(0x00400000), bits)  # unpack -> Python float -> float32 quiets NaNs
    return bits.view(np.float32).reshape(-1, 1)
