import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from triton_client_b200.image_client import PreprocessPlan
from triton_client_b200.utils import cuda_shared_memory as cudashm
from triton_client_b200 import _lib
menu = [(640, 480), (800, 600), (1024, 768), (1280, 720), (1920, 1080), (500, 375), (375, 500), (224, 224), (150, 100)]
n = 8192
pick = np.random.default_rng(1234).choice(len(menu), size=n)
g = torch.Generator(device="cuda").manual_seed(1234)
def run(classes, label):
    sel = [i for i in range(n) if int(pick[i]) in classes]
    pools = {ci: torch.randint(0, 256, (sum(1 for i in sel if pick[i] == ci), menu[ci][1], menu[ci][0], 3), dtype=torch.uint8, device="cuda", generator=g) for ci in classes}
    cur = {ci: 0 for ci in classes}
    sizes, ptrs = [], []
    src_bytes = 0
    for i in sel:
        ci = int(pick[i]); W, H = menu[ci]
        sizes.append((W, H, 3)); ptrs.append(pools[ci][cur[ci]].data_ptr()); cur[ci] += 1
        src_bytes += W * H * 3
    plan = PreprocessPlan(sizes, 1, "FP16", 3, 224, 224, "INCEPTION", device=0)
    m = len(sel)
    nbytes = m * plan.out_bytes_per_image
    region = cudashm.create_shared_memory_region("c5", nbytes, 0)
    for _ in range(3): plan.run(ptrs, region._base_addr, nbytes)
    torch.cuda.synchronize()
    c0 = _lib.kernel_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): plan.run(ptrs, region._base_addr, nbytes)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    algo = src_bytes + nbytes
    print("%-12s n=%5d  %.3f ms  %.0f img/s  %.0f GB/s algorithmic (%.1f%% of 6549)  launches/run %d fused %d" % (
        label, m, ms, m / ms * 1e3, algo / ms / 1e6, algo / ms / 1e6 / 65.49, (_lib.kernel_launch_count() - c0) // 10, plan.info()["fused_images"]))
    plan.close(); cudashm.destroy_shared_memory_region(region)
run(list(range(9)), "cfg5 all")
for ci in range(9): run([ci], "%dx%d" % menu[ci])
