import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from triton_client_b200 import _lib
from triton_client_b200._device import stream_ptr
n, img = 256, 1920 * 1080 * 3
src = np.empty(n * img, dtype=np.uint8); src[:] = 7
dst = torch.empty(n * img, dtype=torch.uint8, device="cuda")
s = stream_ptr(0)
def h2d(chunks):
    per = n // chunks * img
    t0 = time.perf_counter()
    for k in range(chunks):
        _lib.check(_lib.lib.TcbMemcpy(0, dst.data_ptr() + k * per, src.ctypes.data + k * per, per, s, 0))
    torch.cuda.synchronize()
    return time.perf_counter() - t0
for c in (1, 8, 1, 8, 32):
    t = h2d(c); print("H2D staged, %2d chunks: %.1f ms  %.1f GB/s" % (c, t * 1e3, n * img / t / 1e9))
from triton_client_b200.image_client import HostBatchPipeline
from triton_client_b200.utils import cuda_shared_memory as cudashm
pipe = HostBatchPipeline(n, 1920, 1080, 3, 2, "FP32", 3, 224, 224, "VGG", device=0, chunks=8) if True else None
out_bytes = n * pipe.out_bytes_per_image
region = cudashm.create_shared_memory_region("pg", out_bytes, 0)
out_np = np.empty(out_bytes, dtype=np.uint8)
src4 = src.reshape(n, 1080, 1920, 3)
for rep in range(4):
    t0 = time.perf_counter(); pipe.run_pageable(src4, region._base_addr, out_bytes, out_np); torch.cuda.synchronize(); t = time.perf_counter() - t0
    t1 = time.perf_counter(); pipe.run_pageable(src4, region._base_addr, out_bytes, None); torch.cuda.synchronize(); tn = time.perf_counter() - t1
    print("run_pageable with host_out %.1f ms (%.0f img/s), without %.1f ms" % (t * 1e3, n / t, tn * 1e3))
