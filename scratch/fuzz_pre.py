"""One-off fuzz of the preprocess kernels against the C oracle: random geometries, pitches, alignments,
channel counts, layouts and dtypes; a fused/IMMA-eligible share by construction. Prints every mismatch."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import clib
from oracle import pil_bilinear as pb
from triton_client_b200 import _lib
from triton_client_b200.image_client import PreprocessPlan

seed = int(sys.argv[1]) if len(sys.argv) > 1 else 1
trials = int(sys.argv[2]) if len(sys.argv) > 2 else 400
rng = np.random.default_rng(seed)
bad = fused = mma_like = 0
t0 = time.time()
for trial in range(trials):
    kind = rng.integers(0, 4)
    if kind == 0:   # typical downscale
        W, H = int(rng.integers(225, 2200)), int(rng.integers(225, 1300)); w, h = 224, 224
    elif kind == 1:  # odd targets
        W, H = int(rng.integers(16, 1500)), int(rng.integers(9, 900)); w, h = int(rng.integers(1, 400)), int(rng.integers(1, 300))
    elif kind == 2:  # upscale / mixed
        W, H = int(rng.integers(3, 300)), int(rng.integers(3, 300)); w, h = int(rng.integers(8, 513)), int(rng.integers(8, 300))
    else:            # widths around multiples of 16 / 32, heights around multiples of 8
        W = int(rng.choice([16, 32, 48, 64, 96, 640, 1024, 1920])) + int(rng.integers(-2, 3)); H = int(rng.choice([8, 16, 24, 480, 1080])) + int(rng.integers(-2, 3))
        W, H = max(W, 2), max(H, 2); w, h = int(rng.choice([8, 16, 224, 225, 231, 512])), int(rng.choice([7, 8, 224, 226]))
    cin = int(rng.choice([1, 3, 3, 3, 4])); c = 1 if cin == 1 else int(rng.choice([1, 3, 3]))
    fmt = int(rng.choice([1, 2])); dt = str(rng.choice(["FP32", "FP16", "UINT8", "INT16", "FP64"]))
    sc = str(rng.choice(["NONE", "INCEPTION", "VGG"]))
    pitch = W * cin
    if rng.random() < 0.6:
        pitch = (pitch + 15) // 16 * 16 + 16 * int(rng.integers(0, 3))
    else:
        pitch += int(rng.integers(0, 7))
    mis = int(rng.choice([0, 0, 0, 4, 1, 16]))
    n = int(rng.integers(1, 4))
    buf = torch.randint(0, 256, (n * H * pitch + 64,), dtype=torch.uint8, device="cuda")
    host = buf.cpu().numpy()
    ptrs = [buf.data_ptr() + mis + i * H * pitch for i in range(n)]
    path = int(rng.choice([0, 0, 3]))
    try:
        plan = PreprocessPlan([(W, H, cin, pitch)] * n, fmt, dt, c, h, w, sc, device=0, path=path)
        out = torch.zeros(n * plan.out_bytes_per_image, dtype=torch.uint8, device="cuda")
        plan.run(ptrs, out.data_ptr(), out.numel())
    except _lib.TcbError as ex:
        if ex.code == _lib.TCB_ERR_UNSUPPORTED and path == 3:
            continue
        print("ERROR", (W, H, w, h, cin, c, fmt, dt, sc, pitch, mis, path), ex); bad += 1; continue
    torch.cuda.synchronize()
    fused += plan.info()["fused_images"]
    mma_like += int(pitch % 16 == 0 and mis % 16 == 0 and cin == 3 and c == 3)
    got = out.cpu().numpy(); per = plan.out_bytes_per_image
    plan.close()
    lut = pb.scaling_lut(dt, c, sc)
    for i in range(n):
        src = np.lib.stride_tricks.as_strided(host[mis + i * H * pitch:], shape=(H, W, cin), strides=(pitch, cin, 1))
        conv = pb.convert_u8(np.ascontiguousarray(src), c)
        u8 = clib.resize_bilinear_u8(conv, w, h) if (W, H) != (w, h) else conv
        ref = np.empty(u8.shape, dtype=lut.dtype)
        for ch in range(c):
            ref[:, :, ch] = lut[ch][u8[:, :, ch]]
        if fmt == 2:
            ref = np.transpose(ref, (2, 0, 1))
        if got[i * per:(i + 1) * per].tobytes() != np.ascontiguousarray(ref).tobytes():
            print("MISMATCH", (W, H, w, h, cin, c, fmt, dt, sc, pitch, mis, path, n, i)); bad += 1; break
print("seed %d: %d trials, %d bad, %d fused images, %d IMMA-eligible trials, %.0f s" % (seed, trials, bad, fused, mma_like, time.time() - t0))
