import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from triton_client_b200.image_client import PreprocessPlan
from triton_client_b200.utils import cuda_shared_memory as cudashm
def run(n, W, H, path, fmt, dtype, scaling, label):
    g = torch.Generator(device="cuda").manual_seed(1)
    batch = torch.randint(0, 256, (n, H, W, 3), dtype=torch.uint8, device="cuda", generator=g)
    plan = PreprocessPlan([(W, H, 3)] * n, fmt, dtype, 3, 224, 224, scaling, device=0, path=path)
    nbytes = n * plan.out_bytes_per_image
    region = cudashm.create_shared_memory_region("tp", nbytes, 0)
    ptrs = [batch.data_ptr() + i * H * W * 3 for i in range(n)]
    for _ in range(3): plan.run(ptrs, region._base_addr, nbytes)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): plan.run(ptrs, region._base_addr, nbytes)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    algo = n * H * W * 3 + nbytes
    print("%-28s %.3f ms  %.0f GB/s (%.1f%%)" % (label, ms, algo / ms / 1e6, algo / ms / 1e6 / 65.49))
    plan.close(); cudashm.destroy_shared_memory_region(region)
run(256, 1920, 1080, 1, 2, "FP32", "VGG", "cfg3 forced two-pass")
run(908, 224, 224, 0, 1, "FP16", "INCEPTION", "identity 224 NHWC f16")
run(908, 224, 224, 0, 2, "FP32", "VGG", "identity 224 NCHW f32")
run(512, 224, 1080, 0, 2, "FP32", "VGG", "vertical only 224x1080")
