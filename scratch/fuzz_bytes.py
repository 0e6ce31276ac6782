"""One-off fuzz of the BYTES kernels against the C oracle: length distributions (zeros, tiny, uniform,
heavy tail with elements up to 300 KB), payloads that look like headers, element limits, cuts."""
import os, struct, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import bytes_codec as bc
from oracle import clib
from triton_client_b200 import utils as u

seed = int(sys.argv[1]) if len(sys.argv) > 1 else 1
trials = int(sys.argv[2]) if len(sys.argv) > 2 else 200
rng = np.random.default_rng(seed)
bad = 0
t0 = time.time()
for trial in range(trials):
    n = int(rng.choice([1, 2, 7, 100, 1023, 1024, 1025, 5000, 40000, 150000]))
    dist = int(rng.integers(0, 6))
    if dist == 0: lens = np.zeros(n, dtype=np.int64)
    elif dist == 1: lens = rng.integers(0, 4, n)
    elif dist == 2: lens = rng.integers(0, 257, n)
    elif dist == 3: lens = rng.geometric(0.02, n) - 1
    elif dist == 4:
        lens = rng.integers(0, 40, n); k = max(1, n // 500)
        lens[rng.integers(0, n, k)] = rng.integers(600, 300000, k)
    else: lens = rng.integers(500, 9000, min(n, 3000)); n = lens.size
    lens = lens.astype(np.int64)
    P = int(lens.sum())
    if P > 400_000_000: continue
    content = int(rng.integers(0, 3))
    if content == 0: blob = rng.integers(0, 256, P, dtype=np.uint8)
    elif content == 1: blob = np.zeros(P, dtype=np.uint8)
    else:
        blob = np.zeros(P, dtype=np.uint8); blob[::4] = rng.integers(0, 40, (P + 3) // 4, dtype=np.uint8)[: len(blob[::4])]
    want = clib.bytes_serialize(blob, lens)
    d_blob = torch.from_numpy(blob).cuda() if P else torch.empty(1, dtype=torch.uint8, device="cuda")
    ldt = torch.int32 if rng.random() < 0.5 else torch.int64
    d_lens = torch.from_numpy(lens).to(ldt).cuda()
    mis = int(rng.choice([0, 0, 1, 2, 3, 8]))
    outbuf = torch.full((want.size + 64,), 0xA5, dtype=torch.uint8, device="cuda")
    _, total = u.serialize_to_device(d_blob, d_lens, outbuf[mis:mis + want.size], 0, 0)
    got = outbuf.cpu().numpy()
    if total != want.size or not np.array_equal(got[mis:mis + want.size], want) or (got[:mis] != 0xA5).any() or (got[mis + want.size:] != 0xA5).any():
        print("SERIALIZE MISMATCH", (n, dist, content, mis, str(ldt))); bad += 1; continue
    # decode: whole buffer, a limit, a cut
    stream = outbuf[mis:mis + want.size]
    offs, pay = u.deserialize_from_device(stream, want.size, -1, 0, 0)
    if not (np.array_equal(offs.cpu().numpy(), np.concatenate(([0], np.cumsum(lens)))) and np.array_equal(pay.cpu().numpy(), blob)):
        print("DESERIALIZE MISMATCH", (n, dist, content, mis)); bad += 1; continue
    m = int(rng.integers(1, n + 1))
    offs, pay = u.deserialize_from_device(stream, want.size, m, 0, 0)
    wo = np.concatenate(([0], np.cumsum(lens[:m])))
    if not (np.array_equal(offs.cpu().numpy(), wo) and np.array_equal(pay.cpu().numpy(), blob[:wo[-1]])):
        print("LIMIT MISMATCH", (n, dist, content, mis, m)); bad += 1; continue
    if want.size > 8:
        cut = int(rng.integers(1, want.size))
        try:
            w_blob, w_off = bc.deserialize_blob(want[:cut]); ok = True
        except bc.OracleError:
            ok = False
        try:
            offs, pay = u.deserialize_from_device(stream, cut, -1, 0, 0); gok = True
        except struct.error:
            gok = False
        if ok != gok or (ok and not (np.array_equal(offs.cpu().numpy(), w_off) and np.array_equal(pay.cpu().numpy(), w_blob))):
            print("CUT MISMATCH", (n, dist, content, mis, cut, ok, gok)); bad += 1
print("seed %d: %d trials, %d bad, %.0f s" % (seed, trials, bad, time.time() - t0))
