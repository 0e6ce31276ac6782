import os, time, ctypes, mmap
import numpy as np
print("THP enabled:", open("/sys/kernel/mm/transparent_hugepage/enabled").read().strip())
print("THP defrag :", open("/sys/kernel/mm/transparent_hugepage/defrag").read().strip())
print("kernel:", os.uname().release, "numpy", np.__version__)
n = 2112000000
def first_touch(label, arr_fn):
    t0 = time.perf_counter(); a = arr_fn(); t1 = time.perf_counter()
    a[::4096] = 1
    t2 = time.perf_counter()
    print("%-28s alloc %.3f s, touch every page %.3f s (%.1f GB/s)" % (label, t1 - t0, t2 - t1, n / (t2 - t1) / 1e9))
    return a
a = first_touch("np.empty", lambda: np.empty(n, dtype=np.uint8))
del a
libc = ctypes.CDLL("libc.so.6", use_errno=True)
def thp_array():
    m = mmap.mmap(-1, n + (2 << 20), flags=mmap.MAP_PRIVATE | mmap.MAP_ANONYMOUS)
    m.madvise(mmap.MADV_HUGEPAGE)
    return np.frombuffer(m, dtype=np.uint8, count=n)
try:
    a = first_touch("mmap + MADV_HUGEPAGE", thp_array); del a
except Exception as e:
    print("madvise failed", e)
def populate_array():
    m = mmap.mmap(-1, n, flags=mmap.MAP_PRIVATE | mmap.MAP_ANONYMOUS | mmap.MAP_POPULATE)
    return np.frombuffer(m, dtype=np.uint8, count=n)
a = first_touch("mmap MAP_POPULATE", populate_array); del a
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from triton_client_b200.utils import cuda_shared_memory as cudashm
src = np.random.default_rng(1).integers(0, 256, n, dtype=np.uint8)
r = cudashm.create_shared_memory_region("p", n, 0)
cudashm.set_shared_memory_region(r, [src])
for i in range(3):
    t0 = time.perf_counter(); out = cudashm.get_contents_as_numpy(r, np.uint8, [n]); t1 = time.perf_counter()
    print("get_contents_as_numpy: %.1f GB/s" % (n / (t1 - t0) / 1e9))
    del out
from triton_client_b200 import _lib
dst = np.empty(n, dtype=np.uint8); dst[::4096] = 0  # pre-faulted destination
for i in range(3):
    t0 = time.perf_counter()
    _lib.check(_lib.lib.TcbRegionGet(r._c_handle if hasattr(r, "_c_handle") else r._handle, 0, n, dst.ctypes.data)) if False else None
    t1 = time.perf_counter()
# pre-faulted destination through TcbMemcpy (staged D2H)
from triton_client_b200._device import stream_ptr
for i in range(3):
    t0 = time.perf_counter()
    _lib.check(_lib.lib.TcbMemcpy(0, dst.ctypes.data, r._base_addr, n, stream_ptr(0), 1))
    t1 = time.perf_counter()
    print("staged D2H into a pre-faulted array: %.1f GB/s" % (n / (t1 - t0) / 1e9))
print("equal", bool(np.array_equal(dst, src)))
cudashm.destroy_shared_memory_region(r)
