import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from triton_client_b200 import utils as u
N = 1_000_000
rng = np.random.default_rng(42)
lens_h = rng.integers(0, 257, N, dtype=np.int64)
lens = torch.from_numpy(lens_h.astype(np.int32)).cuda()
P = int(lens_h.sum())
blob = torch.randint(0, 256, (P,), dtype=torch.uint8, device="cuda")
out = torch.empty(P + 4 * N + 64, dtype=torch.uint8, device="cuda")
for it in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    u.serialize_to_device(blob, lens, out)
    torch.cuda.synchronize(); print("iter", it, "ms %.3f" % ((time.perf_counter() - t0) * 1e3))
