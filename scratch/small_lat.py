import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from triton_client_b200 import utils as u
from oracle import bytes_codec as bc
def t(fn, n=300):
    for _ in range(20): fn()
    xs = []
    for _ in range(n):
        t0 = time.perf_counter(); fn(); xs.append(time.perf_counter() - t0)
    return np.median(xs) * 1e6
for n in (1, 16, 256, 4096):
    arr = np.empty(n, dtype=np.object_)
    for i in range(n): arr[i] = b"classification:%d:label" % i
    ser = u.serialize_byte_tensor(arr).item()
    assert ser == bc.serialize_elements(arr)
    print("n=%5d serialize %.1f us (python loop %.1f us)   deserialize %.1f us (python loop %.1f us)" % (
        n, t(lambda: u.serialize_byte_tensor(arr)), t(lambda: bc.serialize_elements(arr), 50),
        t(lambda: u.deserialize_bytes_tensor(ser)), t(lambda: bc.deserialize_stream(ser), 50)))
v = np.random.default_rng(0).standard_normal(1000).astype(np.float32)
print("bf16 n=1000: serialize %.1f us  deserialize %.1f us" % (t(lambda: u.serialize_bf16_tensor(v)), t(lambda: u.deserialize_bf16_tensor(u.serialize_bf16_tensor(v).item()))))
