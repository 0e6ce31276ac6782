import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from triton_client_b200 import utils as u, _lib
N = int(os.environ.get("N", 16_000_000))
g = torch.Generator(device="cuda").manual_seed(43)
lens = torch.randint(0, 257, (N,), dtype=torch.int32, device="cuda", generator=g)
P = int(lens.sum(dtype=torch.int64).item())
blob = torch.randint(0, 256, (P,), dtype=torch.uint8, device="cuda", generator=g)
out, total = u.serialize_to_device(blob, lens)
for mode in ("single", "multi"):
    if mode == "multi": os.environ["TCB_DES_NO_SINGLE"] = "1"
    else: os.environ.pop("TCB_DES_NO_SINGLE", None)
    for it in range(4):
        c0 = _lib.kernel_launch_count()
        torch.cuda.synchronize(); t0 = time.perf_counter()
        o, p = u.deserialize_from_device(out, total)
        torch.cuda.synchronize(); t1 = time.perf_counter()
        print(mode, it, "ms %.3f" % ((t1 - t0) * 1e3), "launches", _lib.kernel_launch_count() - c0, "ok", bool(torch.equal(p, blob)))
