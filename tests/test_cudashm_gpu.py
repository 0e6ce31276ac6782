"""GPU tests of the cuda_shared_memory drop-in: the reference's own unit tests
(/root/reference/src/python/library/tests/test_cuda_shared_memory.py:37-164) re-stated against
the new module, plus the error messages and quirks SURVEY.md 8a lists."""
import base64
import struct

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _mods():
    import torch

    import triton_client_b200.utils as utils
    import triton_client_b200.utils.cuda_shared_memory as cudashm

    return torch, utils, cudashm


def test_dlpack_from_gpu():  # DLPackTest.test_from_gpu :42-58
    torch, utils, cudashm = _mods()
    gpu_tensor = torch.ones(4, 4).cuda(0)
    shm_handle = cudashm.create_shared_memory_region("cudashm_data", 64, 0)
    cudashm.set_shared_memory_region_from_dlpack(shm_handle, [gpu_tensor])
    smt = cudashm.as_shared_memory_tensor(shm_handle, "FP32", [4, 4])
    generated = torch.from_dlpack(smt)
    assert torch.allclose(gpu_tensor, generated)
    assert generated.data_ptr() == shm_handle._base_addr  # zero-copy view
    del generated, smt
    cudashm.destroy_shared_memory_region(shm_handle)


def test_dlpack_from_cpu():  # DLPackTest.test_from_cpu :60-80
    torch, utils, cudashm = _mods()
    cpu_tensor = np.ones([4, 4], dtype=np.float32)
    shm_handle = cudashm.create_shared_memory_region("cudashm_data", 64, 0)
    cudashm.set_shared_memory_region_from_dlpack(shm_handle, [cpu_tensor])
    smt = cudashm.as_shared_memory_tensor(shm_handle, "FP32", [4, 4])
    generated = torch.from_dlpack(smt)
    assert np.allclose(cpu_tensor, np.from_dlpack(generated.cpu()))
    del generated, smt
    cudashm.destroy_shared_memory_region(shm_handle)


def test_from_numpy_and_to_numpy():  # NumpyTest.test_from_numpy :88-111, test_to_numpy :113-135
    torch, utils, cudashm = _mods()
    cpu_tensor = np.arange(16, dtype=np.float32).reshape(4, 4)
    shm_handle = cudashm.create_shared_memory_region("cudashm_data", 64, 0)
    cudashm.set_shared_memory_region(shm_handle, [cpu_tensor])
    smt = cudashm.as_shared_memory_tensor(shm_handle, "FP32", [4, 4])
    generated = torch.from_dlpack(smt)
    assert np.allclose(cpu_tensor, np.from_dlpack(generated.cpu()))
    out = cudashm.get_contents_as_numpy(shm_handle, np.float32, [4, 4])
    assert np.array_equal(cpu_tensor, out) and out.flags["OWNDATA"] or out.base is None
    del generated, smt
    cudashm.destroy_shared_memory_region(shm_handle)


def test_numpy_bytes():  # NumpyTest.test_numpy_bytes :137-164
    torch, utils, cudashm = _mods()
    int_tensor = np.arange(start=0, stop=16, dtype=np.int32)
    bytes_tensor = np.array([str(x).encode("utf-8") for x in int_tensor.flatten()], dtype=object)
    bytes_tensor = bytes_tensor.reshape(int_tensor.shape)
    serialized = utils.serialize_byte_tensor(bytes_tensor)
    byte_size = utils.serialized_byte_size(serialized)
    shm_handle = cudashm.create_shared_memory_region("cudashm_data", byte_size, 0)
    cudashm.set_shared_memory_region(shm_handle, [serialized])
    generated = cudashm.get_contents_as_numpy(shm_handle, np.object_, [16])
    assert np.array_equal(bytes_tensor, generated)
    cudashm.destroy_shared_memory_region(shm_handle)


def test_lifecycle_raw_handle_and_region_list():
    torch, utils, cudashm = _mods()
    n0 = len(cudashm.allocated_shared_memory_regions())
    h = cudashm.create_shared_memory_region("r0", 1024, 0)
    assert len(cudashm.allocated_shared_memory_regions()) == n0 + 1
    assert h._byte_size == 1024 and h._device_id == 0 and h._triton_shm_name == "r0"
    assert h._base_addr % 256 == 0
    raw = cudashm.get_raw_handle(h)
    assert isinstance(raw, bytes) and len(raw) == 88 and len(base64.b64decode(raw)) == 64
    assert base64.b64decode(raw) == h._cuda_shm_handle.reserved
    cudashm.destroy_shared_memory_region(h)
    assert len(cudashm.allocated_shared_memory_regions()) == n0
    with pytest.raises(ValueError):
        cudashm.destroy_shared_memory_region(h)


def test_multiple_arrays_pack_back_to_back_and_errors():
    torch, utils, cudashm = _mods()
    a = np.arange(10, dtype=np.int32)
    b = np.arange(6, dtype=np.float64).reshape(2, 3).T  # non-contiguous input
    h = cudashm.create_shared_memory_region("r1", a.nbytes + b.nbytes, 0)
    cudashm.set_shared_memory_region(h, [a, b])
    raw = cudashm.get_contents_as_numpy(h, np.uint8, [a.nbytes + b.nbytes])
    assert raw[: a.nbytes].tobytes() == a.tobytes()
    assert raw[a.nbytes:].tobytes() == np.ascontiguousarray(b).tobytes()
    assert np.array_equal(cudashm.get_contents_as_numpy(h, np.float64, [3, 2], offset=a.nbytes), b)
    with pytest.raises(cudashm.CudaSharedMemoryException, match="must be specified as a numpy array"):
        cudashm.set_shared_memory_region(h, a)
    with pytest.raises(cudashm.CudaSharedMemoryException, match="list/tuple of numpy arrays"):
        cudashm.set_shared_memory_region(h, [a, [1, 2]])
    with pytest.raises(cudashm.CudaSharedMemoryException, match="insufficient"):
        cudashm.get_contents_as_numpy(h, np.float64, [1000])
    with pytest.raises(cudashm.CudaSharedMemoryException, match="unable to set values"):
        cudashm.set_shared_memory_region(h, [np.zeros(1000, dtype=np.float64)])  # bounds-checked here
    assert cudashm.get_contents_as_numpy(h, np.float32, [0]).shape == (0,)
    assert cudashm.get_contents_as_numpy(h, bool, [4]).dtype == np.bool_
    t = torch.arange(12, device="cuda").reshape(3, 4).t()
    with pytest.raises(cudashm.CudaSharedMemoryException, match="not contiguous"):
        cudashm.set_shared_memory_region_from_dlpack(h, [t])
    cudashm.destroy_shared_memory_region(h)


def test_bytes_contents_quirks_and_garbage_tail():
    """get_contents_as_numpy BYTES semantics (cuda_shared_memory/__init__.py:305-323): reads
    prod(shape) elements from offset 0 and ignores whatever follows; shape [] reads one element;
    prod(shape) == 0 reads one element and then fails in reshape; a region too short raises
    struct.error."""
    torch, utils, cudashm = _mods()
    elems = [b"a", b"bc", b"", b"def", b"x" * 700, b"yz"]
    ser = utils.serialize_byte_tensor(np.array(elems, dtype=object))
    n = utils.serialized_byte_size(ser)
    h = cudashm.create_shared_memory_region("r2", n + 20000, 0)
    junk = np.random.default_rng(0).integers(0, 256, n + 20000, dtype=np.uint8)
    cudashm.set_shared_memory_region(h, [junk])          # region pre-filled with garbage
    cudashm.set_shared_memory_region(h, [ser])
    assert list(cudashm.get_contents_as_numpy(h, np.object_, [6])) == elems
    assert cudashm.get_contents_as_numpy(h, np.object_, [2, 3]).shape == (2, 3)
    assert list(cudashm.get_contents_as_numpy(h, np.bytes_, [4])) == elems[:4]
    r = cudashm.get_contents_as_numpy(h, np.object_, [])
    assert r.shape == () and r.item() == b"a"
    with pytest.raises(ValueError):
        cudashm.get_contents_as_numpy(h, np.object_, [0])
    cudashm.destroy_shared_memory_region(h)
    h2 = cudashm.create_shared_memory_region("r3", n, 0)
    cudashm.set_shared_memory_region(h2, [ser])
    with pytest.raises(struct.error):
        cudashm.get_contents_as_numpy(h2, np.object_, [7])  # one more element than the region holds
    cudashm.destroy_shared_memory_region(h2)


@pytest.mark.parametrize("nbytes", [2 * 1024 * 1024, 4 * 1024 * 1024 + 17, 37 * 1024 * 1024 + 5])
def test_large_pageable_copies_go_through_the_staging_ring_and_round_trip(nbytes):
    """set/get of arrays above the staging threshold (staging.cu: pinned ring + host copy threads):
    bytes must survive both directions, at an offset, with neighbours untouched."""
    torch, utils, cudashm = _mods()
    rng = np.random.default_rng(nbytes)
    a = rng.integers(0, 256, nbytes, dtype=np.uint8)
    region = cudashm.create_shared_memory_region("staged", nbytes + 64, 0)
    try:
        guard = np.full(nbytes + 64, 0xEE, dtype=np.uint8)
        cudashm.set_shared_memory_region(region, [guard])
        cudashm.set_shared_memory_region(region, [a], offset=32)
        back = cudashm.get_contents_as_numpy(region, np.uint8, [nbytes + 64])
        assert np.array_equal(back[32:32 + nbytes], a)
        assert (back[:32] == 0xEE).all() and (back[32 + nbytes:] == 0xEE).all()
        # typed view, two arrays back to back (the reference's packing)
        f = rng.standard_normal(nbytes // 8).astype(np.float32)
        cudashm.set_shared_memory_region(region, [f, f[::-1].copy()])
        got = cudashm.get_contents_as_numpy(region, np.float32, [2, f.size])
        assert np.array_equal(got[0], f) and np.array_equal(got[1], f[::-1])
    finally:
        cudashm.destroy_shared_memory_region(region)


def test_large_results_come_in_recycled_page_locked_buffers():
    """get_contents_as_numpy of a large fixed-dtype tensor: right bytes, a new array every call, the
    buffer recycled only after the last view of it died (utils/cuda_shared_memory/_host_pool.py)."""
    import gc

    torch, utils, cudashm = _mods()
    from triton_client_b200.utils.cuda_shared_memory import _host_pool

    n = 3 * (1 << 20) + 123  # > 2 MiB: pooled
    src = np.random.default_rng(3).integers(0, 2**31, n, dtype=np.int32)
    h = cudashm.create_shared_memory_region("pool_data", src.nbytes, 0)
    cudashm.set_shared_memory_region(h, [src])
    try:
        def addr(x):  # not x.ctypes.data inside an assert: pytest's rewritten asserts keep such temporaries alive
            return x.__array_interface__["data"][0]

        a = cudashm.get_contents_as_numpy(h, np.int32, [n])
        assert a.dtype == np.int32 and a.shape == (n,) and a.flags.writeable
        assert np.array_equal(a, src)
        held = _host_pool.pool_stats()["held_bytes"]
        assert held >= src.nbytes
        view = a[1000:2000]
        a_addr = addr(a)
        b = cudashm.get_contents_as_numpy(h, np.int32, [n])
        b_addr = addr(b)
        assert b_addr != a_addr
        assert np.array_equal(b, src)
        a[:10] = -1  # writing one result never shows up in another
        assert np.array_equal(b, src)
        keep = view.copy()
        del a
        gc.collect()
        c = cudashm.get_contents_as_numpy(h, np.int32, [n])  # `view` still holds a's buffer: not reused
        c_addr = addr(c)
        assert c_addr != a_addr and c_addr != b_addr
        assert np.array_equal(view, keep)
        del b, c, view
        gc.collect()
        d = cudashm.get_contents_as_numpy(h, np.int32, [n], offset=0)
        assert np.array_equal(d, src)
        assert addr(d) in (a_addr, b_addr, c_addr)  # a recycled buffer, not a new allocation
        assert _host_pool.pool_stats()["held_bytes"] <= _host_pool._pool.limit
        # offsets and other dtypes
        e = cudashm.get_contents_as_numpy(h, np.uint8, [src.nbytes - 4096], offset=4096)
        assert np.array_equal(e, src.view(np.uint8)[4096:])
    finally:
        cudashm.destroy_shared_memory_region(h)
