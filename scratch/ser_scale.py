import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from triton_client_b200 import utils as u
for N in (250_000, 1_000_000, 4_000_000, 16_000_000):
    g = torch.Generator(device="cuda").manual_seed(43)
    lens = torch.randint(0, 257, (N,), dtype=torch.int32, device="cuda", generator=g)
    P = int(lens.sum(dtype=torch.int64).item())
    blob = torch.randint(0, 256, (P,), dtype=torch.uint8, device="cuda", generator=g)
    out = torch.empty(P + 4 * N + 64, dtype=torch.uint8, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ts = []
    for it in range(6):
        flush.fill_(it)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); u.serialize_to_device(blob, lens, out); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[len(ts) // 2]
    algo = 2 * (P + 4 * N)
    print("N=%9d  %.3f ms  %.0f GB/s algorithmic  (%.1f%% of 6549)" % (N, ms, algo / ms / 1e6, algo / ms / 1e6 / 65.49))
    del blob, out, lens, flush
