import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from triton_client_b200 import utils as u
dev = "cuda:0"
what = sys.argv[1]
if what == "des":
    g = torch.Generator(device=dev).manual_seed(1)
    N = 16_000_000
    lens = torch.randint(0, 257, (N,), dtype=torch.int32, device=dev, generator=g)
    P = int(lens.sum(dtype=torch.int64).item())
    blob = torch.randint(0, 256, (P,), dtype=torch.uint8, device=dev, generator=g)
    stream, tot = u.serialize_to_device(blob, lens)
    for i in range(3):
        print("--- call", i, file=sys.stderr)
        off, pay = u.deserialize_from_device(stream, tot)
    assert torch.equal(pay, blob)
    del blob, stream, off, pay
    n = (256 << 20) // 1024
    lens = torch.full((n,), 1024, dtype=torch.int32, device=dev)
    blob = torch.randint(0, 256, (n * 1024,), dtype=torch.uint8, device=dev)
    stream, tot = u.serialize_to_device(blob, lens)
    print("--- 1KiB elements", file=sys.stderr)
    off, pay = u.deserialize_from_device(stream, tot)
    assert torch.equal(pay, blob)
else:
    n = 64 << 20
    vals = torch.randn(n, device=dev)
    out = torch.empty(2 * n, dtype=torch.uint8, device=dev)
    back = torch.empty(n, dtype=torch.float32, device=dev)
    for name, fn in (("pack", lambda: u.serialize_bf16_to_device(vals, out, 0, 0, sync=False)),
                     ("expand", lambda: u.deserialize_bf16_from_device(out, 2 * n, 0, back, 0, sync=False))):
        ts = []
        for i in range(12):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        t = float(np.median(ts[2:]))
        print(os.environ.get("TCB_BF16_CTAS_PER_SM"), name, "%.1f us" % (t * 1e3), "frac %.3f" % (6 * n / (t * 1e-3) / 1e9 / 6452.2))
