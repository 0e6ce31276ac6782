"""Out-of-bounds writes, checked with guard zones (compute-sanitizer is closed on this GPU pool).

Every kernel family writes into the middle of a poisoned buffer whose capacity is exactly what the
call needs, at 16-byte aligned and at odd offsets; afterwards the bytes in front of and behind the
destination must still hold the poison and the destination must equal the oracle's result. This
catches stray stores (vector stores past a ragged end, a row too many, an unaligned head written
as a full word) wherever the destination is caller memory; the library's own scratch is sized by
the same code that indexes it and is covered by the parity tests only.
"""
import ctypes

import numpy as np
import pytest

from oracle import bytes_codec as bc
from oracle import clib
from oracle import pil_bilinear as pb

pytestmark = pytest.mark.gpu

GUARD = 4096
POISON = 0xA5


class Guarded:
    """`nbytes` of device memory at `misalign` bytes past a 256-byte boundary, poison on both sides."""

    def __init__(self, nbytes, misalign=0):
        import torch

        self.n, self.lo = int(nbytes), GUARD + misalign
        self.buf = torch.full((self.lo + self.n + GUARD,), POISON, dtype=torch.uint8, device="cuda")
        self.view = self.buf[self.lo:self.lo + self.n]
        assert self.buf.data_ptr() % 256 == 0

    def check(self):
        import torch

        torch.cuda.synchronize()
        assert bool((self.buf[:self.lo] == POISON).all()), "bytes in front of the destination were written"
        assert bool((self.buf[self.lo + self.n:] == POISON).all()), "bytes behind the destination were written"
        return self.view.cpu().numpy()


def _oracle_image(src, dtype, scaling, fmt, h, w):
    c = 3
    if src.ndim == 2:
        src = np.repeat(src[:, :, None], 3, axis=2)
    u8 = clib.resize_bilinear_u8(np.ascontiguousarray(src[:, :, :3]), w, h) if src.shape[:2] != (h, w) else src[:, :, :3]
    lut = pb.scaling_lut(dtype, c, scaling)
    out = np.stack([lut[ch][u8[:, :, ch]] for ch in range(c)], axis=-1)
    return np.ascontiguousarray(np.transpose(out, (2, 0, 1))) if fmt == 2 else out


@pytest.mark.parametrize("path", [0, 1, 3])
@pytest.mark.parametrize("fmt,dtype,scaling,h,w,mis", [
    (2, "FP32", "VGG", 224, 224, 0), (1, "FP16", "INCEPTION", 224, 224, 0), (1, "FP16", "INCEPTION", 224, 224, 2),
    (2, "FP32", "NONE", 224, 224, 4), (1, "UINT8", "NONE", 99, 201, 1), (2, "FP64", "INCEPTION", 64, 40, 8),
    (2, "INT16", "NONE", 37, 53, 2)])
def test_preprocess_never_writes_outside_its_destination(path, fmt, dtype, scaling, h, w, mis):
    import torch

    from triton_client_b200 import _lib
    from triton_client_b200.image_client import PreprocessPlan

    rng = np.random.default_rng(11)
    sizes = [(1920, 1080), (640, 480), (500, 375), (224, 224), (150, 100), (w, h), (1280, 720), (33, 700)]
    if path == 3:  # dp4a fused kernel required for every image: leave out what no fused kernel takes
        sizes = [s for s in sizes if s not in ((224, 224), (w, h), (33, 700))]
    imgs = [rng.integers(0, 256, (hh, ww, 3), dtype=np.uint8) for ww, hh in sizes]
    d_imgs = [torch.from_numpy(a).cuda() for a in imgs]
    plan = PreprocessPlan([(ww, hh, 3) for ww, hh in sizes], fmt, dtype, 3, h, w, scaling, device=0, path=path)
    need = len(sizes) * plan.out_bytes_per_image
    g = Guarded(need, mis)
    try:
        plan.run([t.data_ptr() for t in d_imgs], g.view.data_ptr(), need)
    except _lib.TcbError:
        if path == 3:  # still a geometry without a fused kernel: nothing may have been written at all
            assert (g.check() == POISON).all()
            return
        raise
    got = g.check()
    plan.close()
    per = plan.out_bytes_per_image
    for i, a in enumerate(imgs):
        want = _oracle_image(a, dtype, scaling, fmt, h, w)
        assert got[i * per:(i + 1) * per].tobytes() == want.tobytes(), "image %d" % i


@pytest.mark.parametrize("n,mis", [(0, 0), (1, 0), (1, 3), (1000, 0), (1000, 1), (1000, 2), (70001, 3), (70001, 16)])
def test_serialize_and_no_alloc_never_write_outside(n, mis):
    import torch

    from triton_client_b200 import _lib
    from triton_client_b200 import utils as u

    blob, lens = bc.gbytes(n, seed=100 + n) if n else (np.zeros(0, np.uint8), np.zeros(0, np.int64))
    want = clib.bytes_serialize(blob, lens) if n else np.zeros(0, np.uint8)
    d_blob = torch.from_numpy(blob).cuda() if blob.size else torch.empty(1, dtype=torch.uint8, device="cuda")
    d_lens = torch.from_numpy(lens).cuda()
    g = Guarded(want.size, mis)
    _, total = u.serialize_to_device(d_blob, d_lens, g.view, 0, 0)
    assert total == want.size
    assert np.array_equal(g.check(), want)
    g2 = Guarded(want.size, mis)
    tot = ctypes.c_uint64(0)
    _lib.check(_lib.lib.TcbBytesSerializeNoAlloc(0, d_blob.data_ptr(), d_lens.data_ptr(), 8, n, g2.view.data_ptr(),
                                                 want.size, ctypes.byref(tot), None))
    assert tot.value == want.size
    assert np.array_equal(g2.check(), want)


@pytest.mark.parametrize("n,mis,max_len", [(1, 0, 256), (1000, 0, 256), (1000, 8, 256), (50000, 0, 256), (3000, 0, 5000),
                                           (40, 8, 200000)])
def test_deserialize_fill_never_writes_outside_offsets_or_payload(n, mis, max_len):
    import torch

    from triton_client_b200 import _lib

    blob, lens = bc.gbytes(n, seed=7 + n, max_len=max_len)
    ser = clib.bytes_serialize(blob, lens)
    d_ser = torch.from_numpy(np.asarray(ser)).cuda()
    ws_bytes = int(_lib.lib.TcbBytesDeserializeWorkspace(ser.size))
    ws = torch.empty(max(ws_bytes, 1), dtype=torch.uint8, device="cuda")
    cnt, bound = ctypes.c_uint64(0), ctypes.c_uint64(0)
    _lib.check(_lib.lib.TcbBytesDeserializeCount(0, d_ser.data_ptr(), ser.size, -1, ws.data_ptr(), ws_bytes,
                                                 ctypes.byref(cnt), ctypes.byref(bound), None))
    assert cnt.value == n
    g_off = Guarded(8 * (n + 1), mis)            # int64 offsets: 8-byte aligned destinations only
    g_pay = Guarded(blob.size, mis // 8)
    n_out, pb_out = ctypes.c_uint64(0), ctypes.c_uint64(0)
    _lib.check(_lib.lib.TcbBytesDeserializeFill(0, d_ser.data_ptr(), ser.size, -1, g_off.view.data_ptr(), n,
                                                g_pay.view.data_ptr(), blob.size, ws.data_ptr(), ws_bytes,
                                                ctypes.byref(n_out), ctypes.byref(pb_out), None))
    assert n_out.value == n and pb_out.value == blob.size
    offs = g_off.check().view(np.int64)
    assert np.array_equal(offs, np.concatenate(([0], np.cumsum(lens))))
    assert np.array_equal(g_pay.check(), blob)


@pytest.mark.parametrize("n,mis", [(1, 0), (7, 2), (1000, 0), (1001, 2), (65537, 6), (1 << 20, 0)])
def test_bf16_never_writes_outside(n, mis):
    import torch

    from triton_client_b200 import _lib

    rng = np.random.default_rng(n)
    vals = rng.standard_normal(n).astype(np.float32)  # no NaNs: plain truncation (utils/__init__.py:304-313)
    want = (vals.view(np.uint32) >> 16).astype("<u2").tobytes()
    d_vals = torch.from_numpy(vals).cuda()
    g = Guarded(2 * n, mis)
    _lib.check(_lib.lib.TcbBf16Serialize(0, d_vals.data_ptr(), n, g.view.data_ptr(), 2 * n, None))
    packed = g.check()
    assert packed.tobytes() == bytes(want)
    g2 = Guarded(4 * n, (mis // 4) * 4)
    src = torch.from_numpy(packed.copy()).cuda()
    _lib.check(_lib.lib.TcbBf16Deserialize(0, src.data_ptr(), 2 * n, g2.view.data_ptr(), 4 * n, None))
    back = g2.check().view(np.uint32)
    assert np.array_equal(back, (vals.view(np.uint32) >> 16) << 16)


def test_classification_fields_never_write_outside():
    import torch

    from triton_client_b200 import _lib

    elems = [b"0.5:1:a", b"", b"::::", b"x" * 100 + b":" + b"y" * 40, b"1:2"] * 13
    lens = np.array([len(e) for e in elems], dtype=np.int64)
    offs = np.concatenate(([0], np.cumsum(lens))).astype(np.int64)
    payload = torch.from_numpy(np.frombuffer(b"".join(elems), dtype=np.uint8).copy()).cuda()
    d_offs = torch.from_numpy(offs).cuda()
    g = Guarded(32 * len(elems), 8)
    _lib.check(_lib.lib.TcbClassificationFields(0, payload.data_ptr(), d_offs.data_ptr(), len(elems),
                                                g.view.data_ptr(), None))
    got = g.check().view(np.int64).reshape(-1, 4)
    for e, row in zip(elems, got):
        pos = [i for i, ch in enumerate(e) if ch == ord(":")][:3]
        assert row.tolist() == pos + [-1] * (3 - len(pos)) + [len(e)]
