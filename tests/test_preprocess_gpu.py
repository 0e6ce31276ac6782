"""GPU parity tests for the preprocess path: the drop-in ``preprocess`` and the batched
region-writing form, through the C ABI, against (a) the goldens generated from the unmodified
reference and (b) the oracle on the same seeded inputs. Bar: bit-exact (the north_star tolerance
of <= 1e-2 after scaling is met with zero difference)."""
import hashlib
import json
import os

import numpy as np
import pytest

from oracle import clib
from oracle import pil_bilinear as pb
from tests.golden import cases

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")
PRE = json.load(open(os.path.join(GOLD, "preprocess_golden.json")))
SMALL = np.load(os.path.join(GOLD, "preprocess_small.npz"))


def sha(b):
    return hashlib.sha256(b).hexdigest()


def oracle_preprocess(src, c, h, w, fmt, dtype, scaling):
    src = src[:, :, None] if src.ndim == 2 else src
    u8 = clib.resize_bilinear_u8(pb.convert_u8(src, c), w, h)
    lut = pb.scaling_lut(dtype, c, scaling)
    out = np.empty(u8.shape, dtype=lut.dtype)
    for ch in range(c):
        out[:, :, ch] = lut[ch][u8[:, :, ch]]
    return np.ascontiguousarray(np.transpose(out, (2, 0, 1)) if fmt == 2 else out)


@pytest.mark.parametrize("rec", PRE["cases"], ids=[c["name"] for c in PRE["cases"]])
def test_preprocess_dropin_matches_reference_golden(rec):
    from triton_client_b200.image_client import preprocess

    src = cases.make_image(rec)
    img = cases.to_pil(src, rec)
    out = preprocess(img, rec["format"], rec["dtype"], rec["c"], rec["h"], rec["w"], rec["scaling"], "grpc")
    assert str(out.dtype) == rec["out_dtype"]
    assert list(out.shape) == rec["out_shape"]
    outc = np.ascontiguousarray(out)
    if sha(outc.tobytes()) != rec["out_sha256"]:
        ref = oracle_preprocess(src, rec["c"], rec["h"], rec["w"], rec["format"], rec["dtype"], rec["scaling"])
        diff = np.abs(outc.astype(np.float64) - ref.astype(np.float64))
        raise AssertionError("mismatch vs golden: max abs diff %g at %d positions" % (diff.max(), (diff > 0).sum()))
    if rec["name"] in SMALL.files:
        assert np.array_equal(outc, SMALL[rec["name"]])


def _batch_case(sizes, fmt, dtype, c, scaling, h=224, w=224, seed=0, path=0):
    import torch

    from triton_client_b200.image_client import PreprocessPlan
    from triton_client_b200.utils import cuda_shared_memory as cudashm

    rng = np.random.default_rng(seed)
    srcs = [rng.integers(0, 256, (H, W, ch), dtype=np.uint8) for (W, H, ch) in sizes]
    d_imgs = [torch.from_numpy(s).cuda() for s in srcs]
    plan = PreprocessPlan(sizes, fmt, dtype, c, h, w, scaling, device=0, path=path)
    nbytes = plan.n * plan.out_bytes_per_image
    region = cudashm.create_shared_memory_region("pre_batch", nbytes + 64, 0)
    try:
        plan.run([t.data_ptr() for t in d_imgs], region._base_addr + 64, nbytes)
        torch.cuda.synchronize()
        got = cudashm.get_contents_as_numpy(region, np.uint8, [nbytes + 64])[64:]
        got = got.view(plan.out_dtype).reshape(plan.out_shape)
        for i, s in enumerate(srcs):
            ref = oracle_preprocess(s, c, h, w, fmt, dtype, scaling)
            assert np.array_equal(got[i], ref), "image %d %s differs" % (i, sizes[i])
        return plan.info()
    finally:
        plan.close()
        cudashm.destroy_shared_memory_region(region)


def test_batch_mixed_resolution_into_region_nhwc_f16():
    # cfg5's size menu (upscale + identity included), sharded into one region at an offset
    menu = [(640, 480), (800, 600), (1024, 768), (1280, 720), (1920, 1080), (500, 375), (375, 500),
            (224, 224), (150, 100)]
    sizes = [(w, h, 3) for (w, h) in menu] * 2
    _batch_case(sizes, 1, "FP16", 3, "INCEPTION", seed=1234)


def test_batch_same_size_nchw_f32_vgg():
    _batch_case([(1920, 1080, 3)] * 6, 2, "FP32", 3, "VGG", seed=1000)


def test_batch_two_pass_path_forced():
    _batch_case([(640, 480, 3), (1920, 1080, 3), (7, 1000, 3), (224, 300, 3), (300, 224, 3), (224, 224, 3)],
                2, "FP32", 3, "INCEPTION", seed=5, path=1)


def test_batch_gray_and_rgba_sources():
    _batch_case([(640, 480, 4), (320, 240, 1), (641, 479, 3)], 2, "FP32", 1, "VGG", seed=6)
    _batch_case([(640, 480, 4), (320, 240, 1), (641, 479, 3)], 1, "FP16", 3, "NONE", seed=7)


def test_batch_other_targets():
    _batch_case([(640, 480, 3), (1920, 1080, 3)], 2, "FP32", 3, "INCEPTION", h=299, w=299, seed=8)
    _batch_case([(640, 480, 3), (100, 50, 3)], 1, "UINT8", 3, "NONE", h=17, w=31, seed=9)
    _batch_case([(64, 64, 3)], 2, "FP64", 3, "VGG", h=512, w=384, seed=10)


def test_plan_rejects_small_output_and_bad_geometry():
    import torch

    from triton_client_b200 import _lib
    from triton_client_b200.image_client import PreprocessPlan

    plan = PreprocessPlan([(64, 48, 3)], 2, "FP32", 3, 8, 8, "NONE", device=0)
    src = torch.zeros((48, 64, 3), dtype=torch.uint8, device="cuda")
    dst = torch.zeros(8 * 8 * 3, dtype=torch.float32, device="cuda")
    with pytest.raises(_lib.TcbError) as ei:
        plan.run([src.data_ptr()], dst.data_ptr(), 8 * 8 * 3 * 4 - 1)
    assert ei.value.code == _lib.TCB_ERR_OUT_OF_RANGE
    plan.run([src.data_ptr()], dst.data_ptr(), 8 * 8 * 3 * 4)
    torch.cuda.synchronize()
    assert float(dst.abs().sum()) == 0.0
    plan.close()


def test_full_size_properties_cfg3():
    """cfg3 at full size (256 x 1080p -> 224^2 f32 NCHW VGG): checked through properties that do
    not need the oracle on every image: identical images give identical outputs, a constant image
    maps to the constant LUT value, and a strided sample of images equals the oracle."""
    import torch

    from triton_client_b200.image_client import preprocess_batch

    n = 256
    g = torch.Generator(device="cuda").manual_seed(1000)
    batch = torch.randint(0, 256, (n, 1080, 1920, 3), dtype=torch.uint8, device="cuda", generator=g)
    batch[7] = batch[3]
    batch[11] = 200
    out, plan = preprocess_batch(batch, 2, "FP32", 3, 224, 224, "VGG")
    assert out.shape == (n, 3, 224, 224)
    assert torch.equal(out[7], out[3])
    lut = pb.scaling_lut("FP32", 3, "VGG")
    for ch in range(3):
        assert torch.all(out[11, ch] == float(lut[ch, 200]))
    for i in (0, 100, 255):
        ref = oracle_preprocess(batch[i].cpu().numpy(), 3, 224, 224, 2, "FP32", "VGG")
        assert np.array_equal(out[i].cpu().numpy(), ref)
    plan.close()


def test_fused_and_two_pass_paths_agree_on_random_geometries():
    """The fused kernel (path=2) and the general two-pass kernels (path=1) are independent
    implementations; on geometries both accept they must agree bit for bit (and with the oracle
    on a sample)."""
    import torch

    from triton_client_b200.image_client import PreprocessPlan

    from triton_client_b200 import _lib

    rng = np.random.default_rng(99)
    n_both = 0
    for trial in range(16):
        W, H = int(rng.integers(230, 2600)), int(rng.integers(230, 1500))
        w, h = int(rng.choice([224, 299, 160, 64])), int(rng.choice([224, 299, 192, 48]))
        cin = int(rng.choice([1, 3, 4]))
        c = 1 if cin == 1 else int(rng.choice([1, 3]))
        fmt, dt = int(rng.choice([1, 2])), str(rng.choice(["FP32", "FP16", "UINT8", "FP64"]))
        sc = str(rng.choice(["NONE", "INCEPTION", "VGG"]))
        if dt == "UINT8" and sc == "VGG" and c == 1:
            sc = "NONE"
        src = rng.integers(0, 256, (H, W, cin), dtype=np.uint8)
        d_src = torch.from_numpy(src).cuda()
        outs = []
        for path in (1, 2):
            plan = PreprocessPlan([(W, H, cin)], fmt, dt, c, h, w, sc, device=0, path=path)
            out = torch.zeros(plan.out_bytes_per_image, dtype=torch.uint8, device="cuda")
            try:
                plan.run([d_src.data_ptr()], out.data_ptr(), out.numel())
            except _lib.TcbError as ex:   # e.g. more filter taps than the fused kernel's classes cover
                assert path == 2 and ex.code == _lib.TCB_ERR_UNSUPPORTED
                break
            torch.cuda.synchronize()
            assert plan.info()["fused_images"] == (1 if path == 2 else 0)
            outs.append(out.cpu().numpy())
            plan.close()
        if len(outs) < 2:
            continue
        n_both += 1
        assert np.array_equal(outs[0], outs[1]), (W, H, w, h, cin, c, fmt, dt, sc)
        if trial % 3 == 0:
            lutdt = pb.scaling_lut(dt, c, sc).dtype
            ref = oracle_preprocess(src, c, h, w, fmt, dt, sc)
            assert np.array_equal(outs[1].view(lutdt).reshape(ref.shape), ref)
    assert n_both >= 8


def test_fused_path_refuses_what_it_cannot_do():
    import torch

    from triton_client_b200 import _lib
    from triton_client_b200.image_client import PreprocessPlan

    src = torch.zeros((224, 300, 3), dtype=torch.uint8, device="cuda")  # height unchanged: single-axis
    plan = PreprocessPlan([(300, 224, 3)], 2, "FP32", 3, 224, 224, "NONE", device=0, path=2)
    out = torch.zeros(3 * 224 * 224 * 4, dtype=torch.uint8, device="cuda")
    with pytest.raises(_lib.TcbError) as ei:
        plan.run([src.data_ptr()], out.data_ptr(), out.numel())
    assert ei.value.code == _lib.TCB_ERR_UNSUPPORTED and "fused" in ei.value.msg


def test_unaligned_and_pitched_sources():
    """Sources whose rows are not 4/16-byte aligned (odd widths in a packed batch) and sources with
    a row pitch take the byte-wise de-interleave inside the fused kernel; results are unchanged."""
    import torch

    from triton_client_b200.image_client import PreprocessPlan

    rng = np.random.default_rng(5)
    W, H = 641, 479
    packed = rng.integers(0, 256, (3, H, W, 3), dtype=np.uint8)          # images 1, 2 start unaligned
    d_packed = torch.from_numpy(packed).cuda()
    plan = PreprocessPlan([(W, H, 3)] * 3, 1, "FP16", 3, 224, 224, "INCEPTION", device=0)
    out = torch.zeros(3 * plan.out_bytes_per_image, dtype=torch.uint8, device="cuda")
    plan.run([d_packed.data_ptr() + i * H * W * 3 for i in range(3)], out.data_ptr(), out.numel())
    torch.cuda.synchronize()
    got = out.cpu().numpy().view(np.float16).reshape(3, 224, 224, 3)
    for i in range(3):
        assert np.array_equal(got[i], oracle_preprocess(packed[i], 3, 224, 224, 1, "FP16", "INCEPTION"))
    plan.close()
    pitch = 2048                                                          # pitched rows
    wide = rng.integers(0, 256, (480, pitch), dtype=np.uint8)
    d_wide = torch.from_numpy(wide).cuda()
    plan = PreprocessPlan([(640, 480, 3, pitch)], 2, "FP32", 3, 224, 224, "VGG", device=0)
    out = torch.zeros(plan.out_bytes_per_image, dtype=torch.uint8, device="cuda")
    plan.run([d_wide.data_ptr()], out.data_ptr(), out.numel())
    torch.cuda.synchronize()
    ref = oracle_preprocess(wide[:, : 640 * 3].reshape(480, 640, 3), 3, 224, 224, 2, "FP32", "VGG")
    assert np.array_equal(out.cpu().numpy().view(np.float32).reshape(3, 224, 224), ref)
    plan.close()


def test_plan_reused_with_differently_aligned_sources():
    """A plan created with 4-byte aligned sources must stay correct when a later run hands it
    sources that start 1, 2 or 3 bytes off (and back): the row-alignment flag the fused kernel
    reads is refreshed with the pointers."""
    import torch

    from triton_client_b200.image_client import PreprocessPlan

    rng = np.random.default_rng(9)
    W, H = 640, 360
    img = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    ref = oracle_preprocess(img, 3, 224, 224, 2, "FP32", "VGG")
    buf = torch.zeros(H * W * 3 + 64, dtype=torch.uint8, device="cuda")
    plan = PreprocessPlan([(W, H, 3)], 2, "FP32", 3, 224, 224, "VGG", device=0)
    out = torch.zeros(plan.out_bytes_per_image, dtype=torch.uint8, device="cuda")
    for off in (0, 1, 16, 3, 2, 0):
        buf[off: off + H * W * 3].copy_(torch.from_numpy(img).reshape(-1))
        out.zero_()
        plan.run([buf.data_ptr() + off], out.data_ptr(), out.numel())
        torch.cuda.synchronize()
        assert np.array_equal(out.cpu().numpy().view(np.float32).reshape(3, 224, 224), ref), off
    assert plan.info()["fused_images"] == 1
    plan.close()


def test_full_size_properties_cfg5_mixed_resolution():
    """cfg5 at full size: 8192 mixed-resolution images (SURVEY.md 8d size menu, drawn with
    default_rng(1234).choice) -> 224x224 fp16 NHWC INCEPTION in one region. Checked through
    size-independent properties: duplicated sources give identical outputs, constant images map to
    the table constant, identity-size images equal the table applied to the source, and one image
    per size class equals the oracle."""
    import torch

    from triton_client_b200.image_client import PreprocessPlan
    from triton_client_b200.utils import cuda_shared_memory as cudashm

    menu = [(640, 480), (800, 600), (1024, 768), (1280, 720), (1920, 1080), (500, 375), (375, 500), (224, 224), (150, 100)]
    n = 8192
    pick = np.random.default_rng(1234).choice(len(menu), size=n)
    g = torch.Generator(device="cuda").manual_seed(1234)
    pools, cursor = [], [0] * len(menu)
    for ci, (W, H) in enumerate(menu):
        cnt = int((pick == ci).sum())
        pools.append(torch.randint(0, 256, (max(cnt, 1), H, W, 3), dtype=torch.uint8, device="cuda", generator=g))
    pools[4][3] = pools[4][1]        # two identical 1080p sources
    pools[0][2] = 255                # constant images
    pools[2][2] = 0
    sizes, ptrs, where = [], [], []
    for i in range(n):
        ci = int(pick[i])
        W, H = menu[ci]
        sizes.append((W, H, 3))
        ptrs.append(pools[ci][cursor[ci]].data_ptr())
        where.append((ci, cursor[ci]))
        cursor[ci] += 1
    plan = PreprocessPlan(sizes, 1, "FP16", 3, 224, 224, "INCEPTION", device=0)
    nbytes = n * plan.out_bytes_per_image
    region = cudashm.create_shared_memory_region("cfg5", nbytes, 0)
    try:
        plan.run(ptrs, region._base_addr, nbytes)
        torch.cuda.synchronize()
        out = torch.from_dlpack(cudashm.as_shared_memory_tensor(region, "FP16", [n, 224, 224, 3]))
        idx = {w: i for i, w in enumerate(where)}
        assert torch.equal(out[idx[(4, 3)]], out[idx[(4, 1)]])
        lut = pb.scaling_lut("FP16", 3, "INCEPTION")
        assert torch.all(out[idx[(0, 2)]] == float(lut[0, 255]))
        assert torch.all(out[idx[(2, 2)]] == float(lut[0, 0]))
        for ci in range(len(menu)):            # one image per size class against the oracle
            i = idx[(ci, 0)]
            ref = oracle_preprocess(pools[ci][0].cpu().numpy(), 3, 224, 224, 1, "FP16", "INCEPTION")
            assert np.array_equal(out[i].cpu().numpy(), ref), menu[ci]
        info = plan.info()
        assert info["fused_images"] == int((pick != 7).sum())   # everything but the identity-size class
        del out
    finally:
        plan.close()
        cudashm.destroy_shared_memory_region(region)


@pytest.mark.parametrize("chunks", [1, 3, 8])
def test_host_batch_pipeline_matches_oracle(chunks):
    """Pinned host batch -> chunked H2D | kernel | D2H on two streams: every image must come back
    equal to the oracle, whatever the chunking (11 images do not divide evenly)."""
    import torch

    from triton_client_b200.image_client import HostBatchPipeline
    from triton_client_b200.utils import cuda_shared_memory as cudashm

    n, H, W = 11, 240, 320
    rng = np.random.default_rng(77)
    src = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    h_in = torch.from_numpy(src).pin_memory()
    pipe = HostBatchPipeline(n, W, H, 3, 2, "FP32", 3, 56, 72, "INCEPTION", chunks=chunks)
    nbytes = n * pipe.out_bytes_per_image
    h_out = torch.zeros(nbytes, dtype=torch.uint8).pin_memory()
    region = cudashm.create_shared_memory_region("pipe", nbytes, 0)
    try:
        for _ in range(2):  # a second run reuses plans, streams and events
            h_out.zero_()
            pipe.run(h_in.data_ptr(), region._base_addr, nbytes, h_out.data_ptr())
            torch.cuda.synchronize()
            got = h_out.numpy().view(np.float32).reshape(n, 3, 56, 72)
            for i in range(n):
                assert np.array_equal(got[i], oracle_preprocess(src[i], 3, 56, 72, 2, "FP32", "INCEPTION")), i
        on_dev = cudashm.get_contents_as_numpy(region, np.float32, [n, 3, 56, 72])
        assert np.array_equal(on_dev, got)
        with pytest.raises(ValueError):
            pipe.run(h_in.data_ptr(), region._base_addr, nbytes - 1, None)
    finally:
        pipe.close()
        cudashm.destroy_shared_memory_region(region)


@pytest.mark.parametrize("path", [2, 3], ids=["fused-auto", "fused-dp4a"])
def test_plan_run_with_changed_pitch_and_failed_run_keeps_plan_consistent(path):
    """(1) a run may raise an image's pitch - the staging size must follow and the
    result stay exact; (2) a pitch the launch cannot hold is refused; (3) a run that fails
    validation half-way must not leave stale pointers behind."""
    import ctypes

    import torch

    from triton_client_b200 import _lib

    rng = np.random.default_rng(31)
    shapes = [(480, 640), (300, 592)]  # (H, W): rows of 16-byte multiples, same coefficient class (one launch), different staging sizes
    srcs = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for h, w in shapes]
    lut = pb.scaling_lut("FP32", 3, "INCEPTION")
    lut_c = np.ascontiguousarray(lut)

    def padded(a, pitch):
        h, w, _ = a.shape
        buf = torch.zeros((h, pitch), dtype=torch.uint8, device="cuda")
        buf[:, : w * 3] = torch.from_numpy(a.reshape(h, w * 3)).cuda()
        return buf

    def images(bufs, pitches):
        arr = (_lib.TcbImage * 2)()
        for i, (b, p) in enumerate(zip(bufs, pitches)):
            arr[i] = _lib.TcbImage(b.data_ptr(), shapes[i][1], shapes[i][0], p, 3)
        return arr

    p0 = [640 * 3, 592 * 3]
    bufs0 = [padded(a, p) for a, p in zip(srcs, p0)]
    plan = ctypes.c_void_p()
    _lib.check(_lib.lib.TcbPreprocessPlanCreate(0, images(bufs0, p0), 2, 3, 224, 224, _lib.LAYOUT_NCHW,
                                                lut_c.ctypes.data, 4, path, ctypes.byref(plan)))
    out = torch.zeros((2, 3, 224, 224), dtype=torch.float32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    want = [oracle_preprocess(a, 3, 224, 224, 2, "FP32", "INCEPTION") for a in srcs]

    def run(imgs):
        out.zero_()
        rc = _lib.lib.TcbPreprocessPlanRun(plan, imgs, out.data_ptr(), out.numel() * 4, stream)
        torch.cuda.synchronize()
        return rc

    assert run(None) == 0
    for i in range(2):
        assert np.array_equal(out[i].cpu().numpy(), want[i])
    # (1) the SMALLER image's pitch grows (still below the larger image's): accepted and exact
    p1 = [640 * 3, 592 * 3 + 112]  # still a multiple of 16: both fused kernels take it
    bufs1 = [bufs0[0], padded(srcs[1], p1[1])]
    assert run(images(bufs1, p1)) == 0
    for i in range(2):
        assert np.array_equal(out[i].cpu().numpy(), want[i])
    # odd pitch: rows lose their 4-byte alignment, still exact
    p2 = [640 * 3 + 1, 592 * 3 + 3]
    bufs2 = [padded(a, p) for a, p in zip(srcs, p2)]
    rc = run(images(bufs2, p2))
    if path == 3:  # the dp4a kernel reads rows of any alignment
        assert rc == 0
    else:          # the IMMA kernel (chosen at creation for these 16-byte aligned rows) cannot: refused, nothing changed
        assert rc in (0, _lib.TCB_ERR_UNSUPPORTED)
    if rc == 0:
        for i in range(2):
            assert np.array_equal(out[i].cpu().numpy(), want[i])
    else:
        bufs2 = bufs1
    # (2) a pitch far beyond what the launch's shared memory was sized for is refused
    p3 = [640 * 3 + 1, 65536]
    bufs3 = [bufs2[0], padded(srcs[1], p3[1])]
    assert run(images(bufs3, p3)) == _lib.TCB_ERR_UNSUPPORTED
    # (3) ... and that failed run committed nothing: the plan still runs on the previous sources
    assert run(None) == 0
    for i in range(2):
        assert np.array_equal(out[i].cpu().numpy(), want[i])
    # geometry mismatch on image 1 after a changed pointer on image 0: nothing of it sticks either
    bad = images([bufs0[0], bufs0[1]], p0)
    bad[1].width = 593
    assert run(bad) == _lib.TCB_ERR_INVALID_ARG
    del bufs0  # image 0's "new" pointer from the failed call dies here; the plan must not use it
    torch.cuda.empty_cache()
    assert run(None) == 0
    for i in range(2):
        assert np.array_equal(out[i].cpu().numpy(), want[i])
    _lib.lib.TcbPreprocessPlanDestroy(plan)


def test_large_mixed_batch_every_image_exact():
    """600 images of the cfg5 menu in one plan: exercises the cost-balanced work partition (items
    that start and end in the middle of images, CTAs whose share spans images of different sizes)
    and checks EVERY image against the oracle."""
    import torch

    from triton_client_b200.image_client import preprocess_batch

    menu = [(640, 480), (800, 600), (1024, 768), (1280, 720), (1920, 1080), (500, 375), (375, 500), (224, 224), (150, 100)]
    rng = np.random.default_rng(77)
    idx = rng.choice(len(menu), size=600)
    base = {}
    for k, (w, h) in enumerate(menu):  # a few distinct images per size, reused
        base[k] = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for _ in range(3)]
    pick = rng.integers(0, 3, size=600)
    d_base = {k: [torch.from_numpy(a).cuda() for a in v] for k, v in base.items()}
    imgs = [d_base[int(i)][int(j)] for i, j in zip(idx, pick)]
    for fmt, dt, sc in ((1, "FP16", "INCEPTION"), (2, "FP32", "VGG")):
        out, plan = preprocess_batch(imgs, fmt, dt, 3, 224, 224, sc)
        host = out.cpu().numpy()
        want = {(k, j): oracle_preprocess(base[k][j], 3, 224, 224, fmt, dt, sc) for k in base for j in range(3)}
        for n, (i, j) in enumerate(zip(idx, pick)):
            assert np.array_equal(host[n], want[(int(i), int(j))]), (n, menu[int(i)], fmt, dt)
        plan.close()


MODES = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "preprocess_modes_golden.json")))


@pytest.mark.parametrize("rec", MODES["cases"], ids=[c["name"] for c in MODES["cases"]])
def test_preprocess_dropin_other_source_modes_match_reference_golden(rec):
    """CMYK / LA / palette / YCbCr images through the drop-in: Pillow converts (as in the reference),
    the GPU resamples; the result must be the reference's."""
    from triton_client_b200.image_client import preprocess

    src = cases.make_image(rec)
    img = cases.to_pil(src, rec)
    conv = np.asarray(img.convert("L" if rec["c"] == 1 else "RGB"))
    if sha(np.ascontiguousarray(conv).tobytes()) != rec["converted_sha256"]:
        pytest.skip("this Pillow converts the mode differently from the one the goldens were made with")
    out = np.ascontiguousarray(preprocess(img, rec["format"], rec["dtype"], rec["c"], rec["h"], rec["w"], rec["scaling"], "grpc"))
    assert str(out.dtype) == rec["out_dtype"] and list(out.shape) == rec["out_shape"]
    assert sha(out.tobytes()) == rec["out_sha256"]


def test_plans_created_run_and_destroyed_from_several_threads():
    """ctypes drops the GIL during calls: plans of different geometry are created, run many times and
    destroyed by four threads at once (the fused-plan registry used to be an unlocked
    process-global map; the fused state now lives in the plan). Every result must equal the oracle."""
    import threading

    import torch

    from triton_client_b200.image_client import PreprocessPlan

    geoms = [((1920, 1080), 2, "FP32", "VGG"), ((640, 480), 1, "FP16", "INCEPTION"),
             ((500, 375), 2, "FP32", "NONE"), ((1280, 720), 1, "UINT8", "NONE")]
    errors = []

    def worker(k):
        try:
            (W, H), fmt, dtype, scaling = geoms[k]
            rng = np.random.default_rng(8 + k)
            imgs = [rng.integers(0, 256, (H, W, 3), dtype=np.uint8) for _ in range(3)]
            want = [oracle_preprocess(a, 3, 224, 224, fmt, dtype, scaling) for a in imgs]
            d_imgs = [torch.from_numpy(a).cuda() for a in imgs]
            stream = torch.cuda.Stream()
            for rep in range(6):
                plan = PreprocessPlan([(W, H, 3)] * 3, fmt, dtype, 3, 224, 224, scaling, device=0)
                out = torch.empty(3 * plan.out_bytes_per_image, dtype=torch.uint8, device="cuda")
                for _ in range(5):
                    plan.run([t.data_ptr() for t in d_imgs], out.data_ptr(), out.numel(), stream=stream.cuda_stream)
                stream.synchronize()
                got = out.cpu().numpy()
                per = plan.out_bytes_per_image
                for i in range(3):
                    if got[i * per:(i + 1) * per].tobytes() != np.ascontiguousarray(want[i]).tobytes():
                        errors.append("thread %d rep %d image %d differs" % (k, rep, i))
                plan.close()
        except Exception as ex:  # noqa: BLE001
            errors.append("thread %d: %r" % (k, ex))

    threads = [threading.Thread(target=worker, args=(k,)) for k in range(len(geoms))]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=300)
    assert not errors, errors[:3]
