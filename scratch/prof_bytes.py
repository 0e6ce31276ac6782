import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from triton_client_b200 import utils as u
N = int(os.environ.get("N", 2_000_000))
g = torch.Generator(device="cuda").manual_seed(43)
lens = torch.randint(0, 257, (N,), dtype=torch.int32, device="cuda", generator=g)
P = int(lens.sum(dtype=torch.int64).item())
blob = torch.randint(0, 256, (P,), dtype=torch.uint8, device="cuda", generator=g)
out, total = u.serialize_to_device(blob, lens)
o, p = u.deserialize_from_device(out, total)
torch.cuda.synchronize()
print("ok", bool(torch.equal(p, blob)), P, total)
