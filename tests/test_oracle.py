"""CPU tests: the oracle restatements against the committed goldens (generated from the
unmodified reference by tests/golden/make_golden.py) and the reference's own known-answer
vectors. When /root/reference is mounted the restatements are also compared with the
reference directly."""
import hashlib
import json
import os
import struct

import numpy as np
import pytest

from oracle import bytes_codec as bc
from oracle import clib
from oracle import pil_bilinear as pb
from oracle import reference_loader as rl
from tests.golden import cases

GOLD = os.path.join(os.path.dirname(__file__), "golden")
PRE = json.load(open(os.path.join(GOLD, "preprocess_golden.json")))
BYT = json.load(open(os.path.join(GOLD, "bytes_golden.json")))
SMALL = np.load(os.path.join(GOLD, "preprocess_small.npz"))


def sha(b):
    return hashlib.sha256(b).hexdigest()


def _hwc(src):
    return src[:, :, None] if src.ndim == 2 else src


def oracle_preprocess_c(src, case):
    """C resampler + numpy-built LUT (the same split the CUDA path uses)."""
    sample = pb.convert_u8(_hwc(src), case["c"])
    u8 = clib.resize_bilinear_u8(sample, case["w"], case["h"])
    lut = pb.scaling_lut(case["dtype"], case["c"], case["scaling"])
    out = np.empty(u8.shape, dtype=lut.dtype)
    for ch in range(case["c"]):
        out[:, :, ch] = lut[ch][u8[:, :, ch]]
    if case["format"] == pb.FORMAT_NCHW:
        out = np.transpose(out, (2, 0, 1))
    return u8, np.ascontiguousarray(out)


@pytest.mark.parametrize("rec", PRE["cases"], ids=[c["name"] for c in PRE["cases"]])
def test_c_oracle_matches_reference_golden(rec):
    src = cases.make_image(rec)
    assert sha(np.ascontiguousarray(src).tobytes()) == rec["src_sha256"]  # pins the RNG stream
    u8, out = oracle_preprocess_c(src, rec)
    assert sha(np.ascontiguousarray(u8 if u8.shape[2] > 1 else u8[:, :, 0]).tobytes()) == rec["u8_sha256"]
    assert str(out.dtype) == rec["out_dtype"]
    assert list(out.shape) == rec["out_shape"]
    assert sha(out.tobytes()) == rec["out_sha256"]
    if rec["name"] in SMALL.files:
        assert np.array_equal(out, SMALL[rec["name"]])


MODES = json.load(open(os.path.join(GOLD, "preprocess_modes_golden.json")))


@pytest.mark.parametrize("rec", MODES["cases"], ids=[c["name"] for c in MODES["cases"]])
def test_modes_converted_by_pillow_then_oracle_matches_reference_golden(rec):
    """CMYK / LA / P / YCbCr sources: the reference lets Pillow convert them to L or RGB before the
    resize (image_client.py:161-164) and so does the product; from there on it is the oracle's path."""
    src = cases.make_image(rec)
    assert sha(np.ascontiguousarray(src).tobytes()) == rec["src_sha256"]
    conv = np.asarray(cases.to_pil(src, rec).convert("L" if rec["c"] == 1 else "RGB"))
    if sha(np.ascontiguousarray(conv).tobytes()) != rec["converted_sha256"]:
        pytest.skip("this Pillow converts the mode differently from the one the goldens were made with")
    u8, out = oracle_preprocess_c(conv, rec)
    assert str(out.dtype) == rec["out_dtype"] and list(out.shape) == rec["out_shape"]
    assert sha(out.tobytes()) == rec["out_sha256"]


_NUMPY_CASES = [c for c in PRE["cases"] if c["W"] * c["H"] <= 1280 * 720]


@pytest.mark.parametrize("rec", _NUMPY_CASES, ids=[c["name"] for c in _NUMPY_CASES])
def test_numpy_restatement_matches_reference_golden(rec):
    src = cases.make_image(rec)
    out = np.ascontiguousarray(
        pb.preprocess_restated(src, rec["format"], rec["dtype"], rec["c"], rec["h"], rec["w"], rec["scaling"]))
    assert str(out.dtype) == rec["out_dtype"]
    assert sha(out.tobytes()) == rec["out_sha256"]


def test_coeffs_c_equals_numpy():
    for i, o in [(640, 224), (480, 224), (1920, 224), (1080, 224), (3840, 224), (100, 224), (224, 224),
                 (7, 224), (1000, 224), (1, 5), (5, 1), (299, 224), (225, 224)]:
        b1, k1, ks1 = pb.precompute_coeffs(i, o)
        b2, k2, ks2 = clib.precompute_coeffs(i, o)
        assert ks1 == ks2 and np.array_equal(b1, b2) and np.array_equal(k1, k2), (i, o)
        n = b1[:, 1]
        assert (k1.sum(axis=1) > 0).all() and (n >= 1).all() and (b1[:, 0] + n <= i).all()


def test_tap_counts_of_survey():
    # ksize 7 / 7 / 19 / 11 / 37 / 21 and 3 for upscales
    for (i, o), ks in {(640, 224): 7, (480, 224): 7, (1920, 224): 19, (1080, 224): 11,
                       (3840, 224): 37, (2160, 224): 21, (150, 224): 3}.items():
        assert clib.lib().orc_ksize(i, o) == ks


def test_pil_port_matches_golden_when_pillow_matches():
    import PIL

    if PIL.__version__ != PRE["env"]["pillow"]:
        pytest.skip("installed Pillow differs from the one that generated the goldens")
    for rec in PRE["cases"][:8]:
        src = cases.make_image(rec)
        out = np.ascontiguousarray(pb.preprocess_pil(cases.to_pil(src, rec), rec["format"], rec["dtype"],
                                                     rec["c"], rec["h"], rec["w"], rec["scaling"]))
        assert sha(out.tobytes()) == rec["out_sha256"]


# ---- BYTES --------------------------------------------------------------------------------------


def test_bytes_known_answer_cc_client_test():
    # /root/reference/src/c++/tests/cc_client_test.cc:1929-1953, 2140-2153
    wire = bytes([2, 0, 0, 0, 1, 127, 2, 0, 0, 0, 2, 127])
    arr = np.array([b"\x01\x7f", b"\x02\x7f"], dtype=np.object_)
    assert bc.serialize_elements(arr) == wire
    assert bc.deserialize_stream(wire) == [b"\x01\x7f", b"\x02\x7f"]
    blob, off = bc.deserialize_blob(np.frombuffer(wire, dtype=np.uint8))
    assert blob.tobytes() == b"\x01\x7f\x02\x7f" and off.tolist() == [0, 2, 4]
    blob2, off2 = clib.bytes_deserialize(np.frombuffer(wire, dtype=np.uint8))
    assert blob2.tobytes() == blob.tobytes() and off2.tolist() == off.tolist()


def test_bytes_reference_test_numpy_bytes_vector():
    # /root/reference/src/python/library/tests/test_cuda_shared_memory.py:137-164
    arr = np.array([str(x).encode("utf-8") for x in range(16)], dtype=np.object_)
    ser = bc.serialize_elements(arr)
    assert bc.serialized_byte_size(np.asarray(ser, dtype=np.object_)) == len(ser) == 16 * 4 + 10 + 12
    back = np.array(bc.deserialize_stream(ser, max_elements=16), dtype=np.object_)
    assert np.array_equal(back, arr)


@pytest.mark.parametrize("rec", BYT["objects"], ids=[o["name"] for o in BYT["objects"]])
def test_bytes_object_cases(rec):
    arr = cases.object_cases()[rec["name"]]
    assert bc.serialize_elements(arr).hex() == rec["serialized_hex"]
    if arr.size:
        elems = bc.element_bytes(arr)
        blob = np.frombuffer(b"".join(elems), dtype=np.uint8)
        lens = np.array([len(e) for e in elems], dtype=np.int64)
        assert bc.serialize_blob(blob, lens).tobytes().hex() == rec["serialized_hex"]
        assert clib.bytes_serialize(blob, lens).tobytes().hex() == rec["serialized_hex"]
        assert bc.deserialize_stream(bytes.fromhex(rec["serialized_hex"])) == elems


@pytest.mark.parametrize("rec", BYT["gbytes"], ids=[str(g["n"]) for g in BYT["gbytes"]])
def test_gbytes_golden(rec):
    blob, lens = cases.gbytes(rec["n"])
    assert sha(blob.tobytes()) == rec["blob_sha256"]
    ser = clib.bytes_serialize(blob, lens)
    assert ser.shape[0] == rec["serialized_bytes"]
    assert sha(ser.tobytes()) == rec["serialized_sha256"]
    b2, off = clib.bytes_deserialize(ser)
    assert np.array_equal(b2, blob) and np.array_equal(np.diff(off), lens)
    if rec["n"] <= 1000:
        assert bc.serialize_blob(blob, lens).tobytes() == ser.tobytes()
        assert bc.serialize_elements(cases.to_object_array(blob, lens)) == ser.tobytes()
        b3, off3 = bc.deserialize_blob(ser)
        assert np.array_equal(b3, blob) and np.array_equal(off3, off)


def test_bytes_truncation_errors():
    good = bc.serialize_elements(np.array([b"abc", b"de"], dtype=np.object_))
    for cut in (1, 2, 3, 5, 6, len(good) - 1):
        with pytest.raises(struct.error):
            bc.deserialize_stream(good[:cut])
        with pytest.raises(bc.OracleError):
            bc.deserialize_blob(np.frombuffer(good[:cut], dtype=np.uint8))
        with pytest.raises(ValueError):
            clib.bytes_deserialize(np.frombuffer(good[:cut], dtype=np.uint8))
    assert bc.deserialize_stream(b"") == []


def test_cudashm_element_count_quirk():
    # cuda_shared_memory/__init__.py:312 - probed: [] and prod==0 both read one element
    assert bc.cudashm_element_count([4]) == 4
    assert bc.cudashm_element_count([2, 2]) == 4
    assert bc.cudashm_element_count([]) == 1
    assert bc.cudashm_element_count([0]) == 1


# ---- direct comparison with the unmodified reference (build container only) ------------------


@pytest.mark.skipif(not rl.available(), reason="/root/reference not mounted")
def test_restatements_equal_reference_live():
    utils = rl.reference_utils()
    pre = rl.reference_preprocess()
    rng = np.random.default_rng(7)
    for _ in range(6):
        W, H = int(rng.integers(1, 700)), int(rng.integers(1, 500))
        w, h = int(rng.integers(1, 300)), int(rng.integers(1, 300))
        c = int(rng.choice([1, 3]))
        fmt = int(rng.choice([1, 2]))
        dt = str(rng.choice(["FP32", "FP16", "UINT8", "INT16"]))
        sc = str(rng.choice(["NONE", "INCEPTION", "VGG"]))
        src = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
        from PIL import Image

        ref = np.ascontiguousarray(pre(Image.fromarray(src, "RGB"), fmt, dt, c, h, w, sc, "http"))
        case = {"c": c, "w": w, "h": h, "dtype": dt, "scaling": sc, "format": fmt}
        _, mine = oracle_preprocess_c(src, case)
        assert mine.dtype == ref.dtype and np.array_equal(mine, ref), (W, H, w, h, c, fmt, dt, sc)
    for name, arr in cases.object_cases().items():
        r = utils.serialize_byte_tensor(arr)
        ser = r.item() if r.size else b""
        assert bc.serialize_elements(arr) == ser, name
        assert list(utils.deserialize_bytes_tensor(ser)) == bc.deserialize_stream(ser)


# ---- BF16 codec: oracle restatements against goldens from the unmodified reference ----------------
BF = json.load(open(os.path.join(GOLD, "bf16_golden.json")))


@pytest.mark.parametrize("rec", BF["cases"], ids=[c["name"] for c in BF["cases"]])
def test_bf16_oracle_matches_reference_golden(rec):
    arr = cases.bf16_cases()[rec["name"]]
    assert arr.size == rec["n"]
    ser = bc.bf16_serialize(arr)
    assert sha(ser) == rec["serialized_sha256"]
    if arr.size <= 256:
        assert bc.bf16_serialize_elements(arr) == ser
    if rec["serialized_hex"] is not None:
        assert ser.hex() == rec["serialized_hex"]
    back = bc.bf16_deserialize(ser)
    assert list(back.shape) == rec["back_shape"] and str(back.dtype) == rec["back_dtype"]
    assert sha(np.ascontiguousarray(back).tobytes()) == rec["back_sha256"]


def test_bf16_oracle_edges():
    assert bc.bf16_deserialize(b"").shape == (0,)
    with pytest.raises(struct.error):
        bc.bf16_deserialize(b"abc")
    with pytest.raises(bc.OracleError):
        bc.bf16_serialize_elements(np.zeros(3, dtype=np.float64))


@pytest.mark.skipif(not rl.available(), reason="/root/reference not mounted")
def test_bf16_oracle_matches_reference_live():
    u = rl.reference_utils()
    for name, arr in cases.bf16_cases().items():
        ser = u.serialize_bf16_tensor(arr).item()
        assert bc.bf16_serialize(arr) == ser, name
        ref_back = u.deserialize_bf16_tensor(ser)
        got = bc.bf16_deserialize(ser)
        assert got.shape == ref_back.shape and got.tobytes() == ref_back.tobytes(), name
    with pytest.raises(struct.error) as e_ref:
        u.deserialize_bf16_tensor(b"abcde")
    with pytest.raises(struct.error) as e_or:
        bc.bf16_deserialize(b"abcde")
    assert str(e_ref.value) == str(e_or.value)
