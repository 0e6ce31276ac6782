import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from PIL import Image
from triton_client_b200 import _lib, image_client as ic
from triton_client_b200._device import stream_ptr, to_host, current_device, device_buffer
img = Image.fromarray(np.random.default_rng(0).integers(0, 256, (480, 640, 3), dtype=np.uint8), "RGB")
for _ in range(20): ic.preprocess(img, 2, "FP32", 3, 224, 224, "INCEPTION")
def t(fn, n=200):
    xs = []
    for _ in range(n):
        t0 = time.perf_counter(); fn(); xs.append(time.perf_counter() - t0)
    return np.median(xs) * 1e6
print("preprocess()            %.1f us" % t(lambda: ic.preprocess(img, 2, "FP32", 3, 224, 224, "INCEPTION")))
print("_pil_to_array           %.1f us" % t(lambda: ic._pil_to_array(img, 3)))
src = ic._pil_to_array(img, 3)
dev = current_device(); s = stream_ptr(dev)
d_src = device_buffer(src.size, dev)
print("H2D pageable 921 KB     %.1f us" % t(lambda: _lib.lib.TcbMemcpy(dev, d_src.data_ptr(), src.ctypes.data, src.size, s, 1)))
pin = torch.empty(src.size, dtype=torch.uint8, pin_memory=True)
print("copy to pinned          %.1f us" % t(lambda: np.copyto(pin.numpy(), src.reshape(-1))))
print("H2D pinned 921 KB       %.1f us" % t(lambda: _lib.lib.TcbMemcpy(dev, d_src.data_ptr(), pin.data_ptr(), src.size, s, 1)))
key = [k for k in ic._plan_cache][0]
plan, a, b = ic._plan_cache[key][:3]
def run():
    plan.run([a.data_ptr()], b.data_ptr(), plan.out_bytes_per_image, stream=s); _lib.lib.TcbStreamSync(dev, s)
print("plan.run + sync         %.1f us" % t(run))
pout = torch.empty(plan.out_bytes_per_image, dtype=torch.uint8, pin_memory=True)
print("D2H pinned 602 KB       %.1f us" % t(lambda: _lib.lib.TcbMemcpy(dev, pout.data_ptr(), b.data_ptr(), plan.out_bytes_per_image, s, 1)))
print("to_host pageable        %.1f us" % t(lambda: to_host(b.data_ptr(), plan.out_bytes_per_image, dev)))
print("tobytes()               %.1f us" % t(lambda: img.tobytes()))
