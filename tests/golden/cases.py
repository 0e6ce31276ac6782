"""Seeded input recipes shared by make_golden.py (which runs the reference) and the tests (which
do not need it). Recipes follow SURVEY.md 8c/8d."""
import numpy as np

NHWC, NCHW = 1, 2  # ModelInput.Format values used at image_client.py:185


def _case(name, seed, W, H, fmt, dtype, scaling, c=3, w=224, h=224, kind="uniform", mode="RGB"):
    return {"name": name, "seed": seed, "W": W, "H": H, "format": fmt, "dtype": dtype,
            "scaling": scaling, "c": c, "w": w, "h": h, "kind": kind, "mode": mode}


PREPROCESS_CASES = [
    # the six SURVEY.md 8c goldens
    _case("s0_640x480_nchw_f32_inception", 0, 640, 480, NCHW, "FP32", "INCEPTION"),
    _case("s1_1080p_nchw_f32_vgg", 1, 1920, 1080, NCHW, "FP32", "VGG"),
    _case("s2_1080p_nhwc_f16_inception", 2, 1920, 1080, NHWC, "FP16", "INCEPTION"),
    _case("s3_150x100_nhwc_f16_none_upscale", 3, 150, 100, NHWC, "FP16", "NONE"),
    _case("s4_224_identity_nchw_f32_none", 4, 224, 224, NCHW, "FP32", "NONE"),
    _case("s5_500x375_nhwc_f32_vgg", 5, 500, 375, NHWC, "FP32", "VGG"),
    # more geometries (cfg5's size menu and awkward ones)
    _case("s6_800x600", 6, 800, 600, NHWC, "FP16", "INCEPTION"),
    _case("s7_1024x768", 7, 1024, 768, NCHW, "FP32", "VGG"),
    _case("s8_1280x720", 8, 1280, 720, NHWC, "FP16", "INCEPTION"),
    _case("s9_375x500", 9, 375, 500, NCHW, "FP32", "INCEPTION"),
    _case("s10_4k", 10, 3840, 2160, NCHW, "FP32", "VGG"),
    _case("s11_223x225", 11, 223, 225, NHWC, "FP32", "NONE"),
    _case("s12_300x224_w_only", 12, 300, 224, NCHW, "FP32", "NONE"),
    _case("s13_224x300_h_only", 13, 224, 300, NCHW, "FP32", "NONE"),
    _case("s14_1x1", 14, 1, 1, NHWC, "FP32", "INCEPTION"),
    _case("s15_2x3", 15, 2, 3, NCHW, "FP32", "VGG"),
    _case("s16_7x1000_tall_vfirst", 16, 7, 1000, NHWC, "FP32", "NONE"),
    _case("s17_1000x7_wide", 17, 1000, 7, NHWC, "FP32", "NONE"),
    _case("s18_9x1000_vfirst", 18, 9, 1000, NCHW, "FP32", "NONE"),
    _case("s19_10x1000_hfirst", 19, 10, 1000, NCHW, "FP32", "NONE"),
    _case("s20_3x301_vfirst", 20, 3, 301, NCHW, "FP32", "NONE"),
    _case("s21_4000x3000", 21, 4000, 3000, NHWC, "FP16", "VGG"),
    _case("s22_53x37_to_299", 22, 53, 37, NCHW, "FP32", "INCEPTION", w=299, h=299),
    _case("s23_640x480_to_299", 23, 640, 480, NHWC, "FP32", "INCEPTION", w=299, h=299),
    _case("s24_odd_out_17x9", 24, 640, 480, NCHW, "FP32", "VGG", w=17, h=9),
    # grayscale target (c == 1): RGB -> L conversion before the resize
    _case("s25_rgb_to_l_640x480", 25, 640, 480, NCHW, "FP32", "VGG", c=1),
    _case("s26_rgb_to_l_1080p_f16", 26, 1920, 1080, NHWC, "FP16", "INCEPTION", c=1),
    # other source modes handled by convert()
    _case("s27_l_to_rgb", 27, 320, 240, NCHW, "FP32", "INCEPTION", mode="L"),
    _case("s28_rgba_to_rgb", 28, 320, 240, NHWC, "FP32", "VGG", mode="RGBA"),
    _case("s29_l_to_l", 29, 320, 240, NHWC, "FP16", "NONE", c=1, mode="L"),
    _case("s30_rgba_to_l", 30, 123, 77, NCHW, "FP32", "NONE", c=1, mode="RGBA"),
    # structured images exercising rounding edges
    _case("z_all0", 0, 640, 480, NCHW, "FP32", "INCEPTION", kind="zeros"),
    _case("z_all255", 0, 640, 480, NCHW, "FP32", "INCEPTION", kind="ones255"),
    _case("z_checker", 0, 641, 479, NHWC, "FP16", "VGG", kind="checker"),
    _case("z_xgrad", 0, 1920, 1080, NCHW, "FP32", "NONE", kind="xgrad"),
    _case("z_ygrad", 0, 1280, 720, NCHW, "FP32", "NONE", kind="ygrad"),
    # numpy promotion oddities
    _case("d_uint8_inception_f64", 31, 640, 480, NCHW, "UINT8", "INCEPTION"),
    _case("d_uint8_vgg_wrap", 32, 640, 480, NHWC, "UINT8", "VGG"),
    _case("d_int32_none", 33, 320, 240, NCHW, "INT32", "NONE"),
    _case("d_fp64_vgg", 34, 320, 240, NHWC, "FP64", "VGG"),
    _case("d_int8_vgg", 35, 320, 240, NHWC, "INT8", "VGG"),
    _case("d_uint16_inception", 36, 320, 240, NHWC, "UINT16", "INCEPTION"),
    # tiny outputs stored in full (preprocess_small.npz)
    _case("t_small_64x48_to_8x8", 40, 64, 48, NCHW, "FP32", "INCEPTION", w=8, h=8),
    _case("t_small_up_5x4_to_16x12", 41, 5, 4, NHWC, "FP16", "VGG", w=16, h=12),
    _case("t_small_31x17_to_13x11_l", 42, 31, 17, NCHW, "FP32", "VGG", c=1, w=13, h=11),
]


# Source modes the kernels do not take as they are: the reference converts them with Pillow
# (image_client.py:161-164: img.convert("L") / img.convert("RGB")) before the resize, and so does the
# drop-in (image_client._pil_to_array) - file-format glue, not the resampling path. These cases pin
# that the glue is the same.
MODE_CASES = [
    _case("m_cmyk_to_rgb", 50, 320, 240, NCHW, "FP32", "INCEPTION", mode="CMYK"),
    _case("m_la_to_rgb", 51, 333, 211, NHWC, "FP16", "VGG", mode="LA"),
    _case("m_la_to_l", 52, 320, 240, NCHW, "FP32", "NONE", c=1, mode="LA"),
    _case("m_palette_to_rgb", 53, 640, 480, NCHW, "FP32", "VGG", mode="P"),
    _case("m_palette_to_l", 54, 200, 300, NHWC, "UINT8", "NONE", c=1, mode="P"),
    _case("m_ycbcr_to_rgb", 55, 500, 375, NHWC, "FP32", "INCEPTION", mode="YCbCr"),
    _case("m_cmyk_to_l", 56, 150, 100, NCHW, "FP16", "INCEPTION", c=1, mode="CMYK"),
]


def make_image(case):
    """uint8 source array: [H, W, 3] for RGB, [H, W] for L, [H, W, 4] for RGBA (and the raw planes of
    the other modes: CMYK 4, LA 2, P 1 index plane, YCbCr 3)."""
    H, W = case["H"], case["W"]
    nch = {"RGB": 3, "L": 1, "RGBA": 4, "CMYK": 4, "LA": 2, "P": 1, "YCbCr": 3}[case["mode"]]
    kind = case["kind"]
    if kind == "uniform":
        a = np.random.default_rng(case["seed"]).integers(0, 256, (H, W, nch), dtype=np.uint8)
    elif kind == "zeros":
        a = np.zeros((H, W, nch), dtype=np.uint8)
    elif kind == "ones255":
        a = np.full((H, W, nch), 255, dtype=np.uint8)
    elif kind == "checker":
        yy, xx = np.mgrid[0:H, 0:W]
        a = (((yy + xx) & 1) * 255).astype(np.uint8)[:, :, None].repeat(nch, axis=2)
    elif kind == "xgrad":
        a = ((np.arange(W) * 255) // max(W - 1, 1)).astype(np.uint8)[None, :, None].repeat(H, 0).repeat(nch, 2)
    elif kind == "ygrad":
        a = ((np.arange(H) * 255) // max(H - 1, 1)).astype(np.uint8)[:, None, None].repeat(W, 1).repeat(nch, 2)
    else:
        raise ValueError(kind)
    a = np.ascontiguousarray(a)
    if nch == 1:
        a = a[:, :, 0]
    return a


def to_pil(src, case):
    from PIL import Image

    if case["mode"] == "P":  # an index plane + a palette drawn from the case's seed
        img = Image.fromarray(src, "P")
        pal = np.random.default_rng(case["seed"] + 1000).integers(0, 256, 768, dtype=np.uint8)
        img.putpalette(pal.tolist())
        return img
    if case["mode"] in ("CMYK", "LA", "YCbCr"):
        return Image.frombytes(case["mode"], (case["W"], case["H"]), np.ascontiguousarray(src).tobytes())
    return Image.fromarray(src, case["mode"])


# ---- BYTES ------------------------------------------------------------------------------------
GBYTES_N = [16, 1000, 1_000_000]


def gbytes(n, seed=42, max_len=256):
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


def object_cases():
    """Inputs whose per-element normalisation is non-trivial (utils/__init__.py:229-239)."""
    c = {}
    c["bytes_1d"] = np.array([b"a", b"", b"bcd", b"\x00\x01\x00"], dtype=np.object_)
    c["str_elems"] = np.array(["he\x00llo", "wörld", ""], dtype=np.object_)
    c["mixed_objs"] = np.array([b"x", "y", 12, None, 3.5, np.bytes_(b"zz")], dtype=np.object_)
    c["bytes_2d_fortran"] = np.asfortranarray(
        np.array([[b"00", b"01", b"02"], [b"10", b"11", b"12"]], dtype=np.object_))
    c["bytes_view_T"] = np.array([[b"00", b"01", b"02"], [b"10", b"11", b"12"]], dtype=np.object_).T
    c["np_bytes_S4_trailing_nul"] = np.array([b"\x00\x01\x02\x03", b"ab\x00\x00", b"\x00\x00\x00\x00", b"abcd"],
                                             dtype=np.bytes_)
    c["np_bytes_S_2d"] = np.array([[b"a", b"bcd"], [b"", b"ef"]], dtype="S3")
    c["empty_obj"] = np.empty((0,), dtype=np.object_)
    c["empty_2d"] = np.empty((3, 0), dtype=np.object_)
    c["zero_d"] = np.array(b"solo", dtype=np.object_)
    # reference usage vectors
    c["ref_test_numpy_bytes"] = np.array([str(x).encode("utf-8") for x in range(16)], dtype=np.object_)
    c["ref_string_infer_object"] = np.array([["he\x00llo"] * 16], dtype=np.object_)
    c["ref_string_infer_bytes_"] = np.array([[b"\x00\x01\x02\x03"] * 16], dtype=np.bytes_)
    c["ref_cc_client_test"] = np.array([b"\x01\x7f", b"\x02\x7f"], dtype=np.object_)
    return c


# ---- BF16 codec cases ----------------------------------------------------------------------------
BF16_SEED = 77


def bf16_cases():
    """name -> float32 ndarray. Specials (NaN with payload, +-Inf, denormals, -0, max), a
    non-contiguous view, a Fortran-ordered array and a seeded random block of raw bit patterns."""
    rng = np.random.default_rng(BF16_SEED)
    specials = np.array([0x00000000, 0x80000000, 0x7F800000, 0xFF800000, 0x7FC00000, 0x7FC12345, 0xFFFFFFFF,
                         0x00000001, 0x00007FFF, 0x00008000, 0x0000FFFF, 0x00010000, 0x007FFFFF, 0x7F7FFFFF,
                         0x3F800000, 0x3F80FFFF, 0x3F808000, 0x3F807FFF, 0xC0490FDB,
                         0x7F800001, 0xFF800001, 0x7FA00000, 0x7FBFFFFF, 0xFF80FFFF], dtype=np.uint32).view(np.float32)
    raw = rng.integers(0, 1 << 32, 4099, dtype=np.uint64).astype(np.uint32).view(np.float32)
    mat = rng.standard_normal((37, 53)).astype(np.float32)
    return {
        "specials": specials,
        "raw_bits_4099": raw,
        "scalar_0d": np.float32(3.14159).reshape(()),
        "matrix": mat,
        "transposed_view": mat.T,
        "strided_view": mat[::3, 1::2],
        "fortran": np.asfortranarray(mat),
        "single": np.array([1.5], dtype=np.float32),
    }


# ---- HTTP request / response cases -------------------------------------------
def http_request_cases(h):
    """name -> thunk returning (inputs, outputs, kwargs) built with the classes of module `h` (the
    reference's tritonclient.http or triton_client_b200.http). Thunks, because building a BYTES / BF16
    input serialises it (on the GPU in this repo). Integer JSON data only: rapidjson and json print
    some floats differently and the reference's rapidjson is a stand-in here."""

    def fixed():
        a = h.InferInput("INPUT0", [2, 2], "INT32").set_data_from_numpy(np.arange(4, dtype=np.int32).reshape(2, 2))
        b = h.InferInput("INPUT1", [3], "INT64").set_data_from_numpy(np.array([7, -8, 9], dtype=np.int64), binary_data=False)
        return [a, b], None, {}

    def text():
        s = h.InferInput("TEXT", [3], "BYTES").set_data_from_numpy(np.array([b"ab", "he\x00llo", 17], dtype=np.object_))
        t = h.InferInput("TEXTJ", [2], "BYTES").set_data_from_numpy(np.array([b"x", "yz"], dtype=np.object_), binary_data=False)
        o = h.InferRequestedOutput("OUT0")
        oj = h.InferRequestedOutput("OUT1", binary_data=False)
        oc = h.InferRequestedOutput("CLS", class_count=3)
        return [s, t], [o, oj, oc], {"request_id": "42"}

    def bf16():
        rng = np.random.default_rng(5)
        f = h.InferInput("BF", [2, 3], "BF16").set_data_from_numpy(rng.standard_normal((2, 3)).astype(np.float32))
        return [f], [h.InferRequestedOutput("OUT0")], {}

    def shm():
        m = h.InferInput("SHM", [4, 8], "FP32").set_shared_memory("region_a", 128, offset=64)
        om = h.InferRequestedOutput("OUTSHM")
        om.set_shared_memory("region_b", 256)
        return [m], [om], {"sequence_id": 5, "sequence_start": True, "priority": 2, "timeout": 1000,
                           "parameters": {"foo": "bar", "n": 3}}

    def big():
        t = h.InferInput("BIGTEXT", [40, 25], "BYTES").set_data_from_numpy(to_object_array(*gbytes(1000)).reshape(40, 25))
        return [t], None, {"sequence_id": "seq-a", "sequence_end": True}

    def empty():
        e = h.InferInput("EMPTY", [0], "BYTES").set_data_from_numpy(np.array([], dtype=np.object_))
        return [e], None, {}

    return {"fixed_binary_and_json_no_outputs": fixed, "bytes_binary_and_json_with_outputs": text, "bf16_binary": bf16,
            "shared_memory_refs": shm, "gbytes_1000": big, "empty_bytes": empty}


def http_response_body():
    """A response as a server would send it: JSON header + binary sections (FP32, BYTES, BF16, an
    empty tensor) and one JSON-data output. Returns (body, header_length)."""
    import json
    import struct

    fp = np.arange(6, dtype=np.float32).reshape(2, 3)
    strs = [b"alpha", b"", b"\x00\x01\x02", "résumé".encode()]
    bytes_sec = b"".join(struct.pack("<I", len(x)) + x for x in strs)
    bf = (np.array([1.5, -2.0, 3.140625], dtype=np.float32).view(np.uint32) >> 16).astype("<u2").tobytes()
    hdr = {"model_name": "m", "model_version": "1", "id": "42", "outputs": [
        {"name": "OUT_FP32", "datatype": "FP32", "shape": [2, 3], "parameters": {"binary_data_size": fp.nbytes}},
        {"name": "OUT_BYTES", "datatype": "BYTES", "shape": [2, 2], "parameters": {"binary_data_size": len(bytes_sec)}},
        {"name": "OUT_JSON", "datatype": "INT64", "shape": [3], "data": [1, 2, 3]},
        {"name": "OUT_BF16", "datatype": "BF16", "shape": [3], "parameters": {"binary_data_size": len(bf)}},
        {"name": "OUT_EMPTY", "datatype": "INT32", "shape": [0], "parameters": {"binary_data_size": 0}},
    ]}
    hj = json.dumps(hdr, separators=(",", ":")).encode()
    return hj + fp.tobytes() + bytes_sec + bf, len(hj)


def postprocess_cases():
    """Classification outputs as a server's classification extension sends them
    ("score:index:label" per element): name -> (object array, batch_size, supports_batching).
    A case whose value list ends with an element of fewer than three fields raises in the
    reference after the lines before it were printed."""
    def arr(rows):
        a = np.empty((len(rows), len(rows[0])), dtype=np.object_)
        for i, r in enumerate(rows):
            for j, x in enumerate(r):
                a[i, j] = x
        return a

    def arr1(items):
        a = np.empty(len(items), dtype=np.object_)
        for i, x in enumerate(items):
            a[i] = x
        return a

    rng = np.random.default_rng(77)
    big = [[("%.6f:%d:label %d/%d" % (rng.random(), rng.integers(0, 1000), i, j)).encode() for j in range(5)]
           for i in range(64)]
    long_label = b"0.5:3:" + bytes(rng.integers(32, 127, 300, dtype=np.uint8).tolist()).replace(b":", b";")
    return {
        "batch2_top3": (arr([[b"0.87:281:tabby cat", b"0.09:282:tiger cat", b"0.01:285:Egyptian cat"],
                             [b"0.5:1:goldfish", b"0.25:2:shark", b"0.125:3:ray"]]), 2, True),
        "non_batching_top2": (arr1([b"12.5:7:cock", b"11.0:8:hen"]), 1, False),
        "label_with_colons": (arr([[b"0.1:7:a:b:c", b"1:2:3:"]]), 1, True),
        "empty_fields": (arr([[b"::", b":5:", b"x::y"]]), 1, True),
        "latin1_bytes": (arr([[b"\xe9:1:caf\xe9", bytes(range(128, 140)) + b":0:" + bytes(range(250, 256))]]), 1, True),
        "colon_at_word_edges": (arr([[b"0123456789abcdef0123456789abcde:1:2", b"0123456789abcdef0123456789abcdef:1:2",
                                      b"a" * 63 + b":" + b"b" * 32 + b":" + b"c" * 70]]), 1, True),
        "long_label": (arr([[long_label]]), 1, True),
        "batch64_top5": (arr(big), 64, True),
        "too_few_fields_raises": (arr([[b"0.9:1:ok", b"0.1:2"]]), 1, True),
        "batch_mismatch_raises": (arr([[b"0.9:1:ok"]]), 2, True),
    }
