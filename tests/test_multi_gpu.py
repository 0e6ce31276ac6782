"""GPU tests of the sharded path and of what a region is for (another process opens it).

* two devices (skipped below two): every "rank" preprocesses / serialises its shard_range of the work
  into its own region on its own device; the per-rank outputs concatenated in rank order must equal
  the unsharded oracle result - for preprocess and for BYTES.
* one device: the same concatenation property with two ranks' shards run one after the other (so
  the property is exercised on a single-GPU box too).
* a second PROCESS opens base64.b64decode(get_raw_handle(region)) with cudaIpcOpenMemHandle and reads
  back the bytes a kernel of this process wrote (what the server does:
  /root/reference/src/python/library/tritonclient/utils/cuda_shared_memory/__init__.py:152-170,
  grpc/_client.py:1379-1384).
"""
import base64
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle import bytes_codec as bc
from oracle import clib
from oracle import pil_bilinear as pb

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SIZES = [(640, 480), (1920, 1080), (500, 375), (150, 100), (224, 224), (375, 500), (1280, 720), (800, 600),
         (1024, 768), (640, 480), (1920, 1080)]


def _oracle_image(src, dtype, scaling, fmt):
    u8 = clib.resize_bilinear_u8(src, 224, 224)
    lut = pb.scaling_lut(dtype, 3, scaling)
    out = np.stack([lut[ch][u8[:, :, ch]] for ch in range(3)], axis=-1)
    return np.ascontiguousarray(np.transpose(out, (2, 0, 1))) if fmt == 2 else out


def _run_sharded(devices):
    """len(devices) ranks; rank r uses devices[r]. Returns the per-rank host outputs."""
    import torch

    from triton_client_b200 import utils as u
    from triton_client_b200.image_client import preprocess_batch
    from triton_client_b200.sharding import exclusive_offsets, shard_range
    from triton_client_b200.utils import cuda_shared_memory as cudashm

    world = len(devices)
    rng = np.random.default_rng(2024)
    images = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for (w, h) in SIZES]
    blob, lens = bc.gbytes(20011, seed=9)
    starts = np.concatenate(([0], np.cumsum(lens)))
    pre_parts, ser_parts = [], []
    for r, dev in enumerate(devices):
        torch.cuda.set_device(dev)
        lo, hi = shard_range(len(images), r, world)
        d_imgs = [torch.from_numpy(a).to("cuda:%d" % dev) for a in images[lo:hi]]
        per = 224 * 224 * 3 * 2
        region = cudashm.create_shared_memory_region("shard_pre_%d" % r, (hi - lo) * per, dev)
        _, plan = preprocess_batch(d_imgs, 1, "FP16", 3, 224, 224, "INCEPTION", out=region, device=dev)
        pre_parts.append(cudashm.get_contents_as_numpy(region, np.float16, [hi - lo, 224, 224, 3]))
        plan.close()
        cudashm.destroy_shared_memory_region(region)
        # BYTES: this rank's element range into its own region
        elo, ehi = shard_range(len(lens), r, world)
        my_blob, my_lens = blob[starts[elo]:starts[ehi]], lens[elo:ehi]
        need = int(my_blob.size) + 4 * int(my_lens.size)
        region = cudashm.create_shared_memory_region("shard_ser_%d" % r, need, dev)
        d_blob = torch.from_numpy(my_blob).to("cuda:%d" % dev) if my_blob.size else torch.empty(1, dtype=torch.uint8, device="cuda:%d" % dev)
        _, total = u.serialize_to_device(d_blob, torch.from_numpy(my_lens).to("cuda:%d" % dev), region, 0, dev)
        assert total == need
        ser_parts.append(cudashm.get_contents_as_numpy(region, np.uint8, [need]))
        cudashm.destroy_shared_memory_region(region)
    torch.cuda.set_device(devices[0])
    # unsharded oracle results
    want_pre = np.stack([_oracle_image(a, "FP16", "INCEPTION", 1) for a in images])
    want_ser = clib.bytes_serialize(blob, lens)
    got_pre = np.concatenate(pre_parts)
    assert got_pre.shape == want_pre.shape
    assert np.array_equal(got_pre.view(np.uint16), want_pre.view(np.uint16))
    got_ser = np.concatenate(ser_parts)
    assert np.array_equal(got_ser, want_ser)
    # base offsets of the per-rank pieces inside the logical tensor: an exclusive sum, no collective
    offs, tot = exclusive_offsets([p.size for p in ser_parts])
    assert tot == want_ser.size
    for r in range(world):
        assert np.array_equal(want_ser[offs[r]:offs[r] + ser_parts[r].size], ser_parts[r])


def test_two_ranks_one_device_concatenate_to_unsharded_oracle():
    _run_sharded([0, 0])


def test_three_ranks_one_device_uneven_shards():
    _run_sharded([0, 0, 0])


def test_two_devices_concatenate_to_unsharded_oracle():
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    _run_sharded([0, 1])


_CHILD = r'''
import base64, ctypes, sys
import numpy as np
handle_b64, nbytes, dev = sys.argv[1].encode(), int(sys.argv[2]), int(sys.argv[3])
rt = None
for name in ("/usr/local/cuda/lib64/libcudart.so", "libcudart.so", "libcudart.so.12", "libcudart.so.13"):
    try:
        rt = ctypes.CDLL(name); break
    except OSError:
        pass
if rt is None:
    import glob, torch, os
    cands = glob.glob(os.path.join(os.path.dirname(torch.__file__), "..", "nvidia", "cuda_runtime", "lib", "libcudart.so*"))
    cands += glob.glob("/usr/local/cuda/lib64/libcudart.so*")
    rt = ctypes.CDLL(sorted(cands)[0])
class Handle(ctypes.Structure):
    _fields_ = [("reserved", ctypes.c_char * 64)]
raw = base64.b64decode(handle_b64)
assert len(raw) == 64
h = Handle.from_buffer_copy(raw)
assert rt.cudaSetDevice(dev) == 0
ptr = ctypes.c_void_p()
rc = rt.cudaIpcOpenMemHandle(ctypes.byref(ptr), h, 1)  # cudaIpcMemLazyEnablePeerAccess
assert rc == 0, "cudaIpcOpenMemHandle failed: %d" % rc
out = np.empty(nbytes, dtype=np.uint8)
rc = rt.cudaMemcpy(ctypes.c_void_p(out.ctypes.data), ptr, ctypes.c_size_t(nbytes), 2)  # DeviceToHost
assert rc == 0, "cudaMemcpy failed: %d" % rc
rt.cudaIpcCloseMemHandle(ptr)
sys.stdout.buffer.write(out.tobytes())
'''


def test_second_process_opens_the_ipc_handle_and_reads_kernel_output():
    import torch

    from triton_client_b200 import utils as u
    from triton_client_b200.utils import cuda_shared_memory as cudashm

    blob, lens = bc.gbytes(3000, seed=5)
    want = clib.bytes_serialize(blob, lens)
    region = cudashm.create_shared_memory_region("ipc_region", want.size, 0)
    try:
        _, total = u.serialize_to_device(torch.from_numpy(blob).cuda(), torch.from_numpy(lens).cuda(), region, 0, 0)
        assert total == want.size
        handle = cudashm.get_raw_handle(region)
        assert len(base64.b64decode(handle)) == 64
        env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
        p = subprocess.run([sys.executable, "-c", _CHILD, handle.decode(), str(want.size), "0"], capture_output=True,
                           timeout=180, env=env)
        assert p.returncode == 0, p.stderr.decode()[-800:]
        got = np.frombuffer(p.stdout, dtype=np.uint8)
        assert got.size == want.size and np.array_equal(got, want)
    finally:
        cudashm.destroy_shared_memory_region(region)
