"""System shared-memory drop-in: the reference's own unit-test shapes
(/root/reference/src/python/library/tests/test_shared_memory.py) restated against
triton_client_b200.utils.shared_memory, plus the device bridge and the GPU BYTES decode."""
import os
import struct

import numpy as np
import pytest

import triton_client_b200.utils.shared_memory as shm
from oracle import bytes_codec as bc
from tests.golden import cases


def _key(name):
    return "/tcb_%s_%d" % (name, os.getpid())


def test_lifecycle_set_get_offsets_and_errors():
    h = shm.create_shared_memory_region("dummy_data", _key("life"), 64)
    try:
        assert shm.mapped_shared_memory_regions() == [_key("life")]
        a = np.arange(4, dtype=np.int32)
        b = np.array([1.5, 2.5], dtype=np.float32)
        shm.set_shared_memory_region(h, [a, b])  # back to back from offset 0
        assert np.array_equal(shm.get_contents_as_numpy(h, np.int32, [2, 2]), a.reshape(2, 2))
        assert np.array_equal(shm.get_contents_as_numpy(h, np.float32, [2], offset=16), b)
        shm.set_shared_memory_region(h, [b], offset=32)
        assert np.array_equal(shm.get_contents_as_numpy(h, np.float32, [2], offset=32), b)
        # a second mapping of the same key sees the bytes (what the server does with the key)
        h2 = shm.create_shared_memory_region("dummy_data2", _key("life"), 64)
        assert np.array_equal(shm.get_contents_as_numpy(h2, np.int32, [4]), a)
        shm.mapped_shm_regions.remove(_key("life"))  # the second handle shares the key: unmap without unlinking twice
        with pytest.raises(shm.SharedMemoryException, match="insufficient to provide numpy array"):
            shm.get_contents_as_numpy(h, np.int64, [9])
        with pytest.raises(shm.SharedMemoryException, match="exceeds region of 64 bytes"):
            shm.set_shared_memory_region(h, [np.zeros(100)])
        with pytest.raises(shm.SharedMemoryException, match="input_values must be specified as a list/tuple of numpy arrays"):
            shm.set_shared_memory_region(h, a)
        with pytest.raises(shm.SharedMemoryException, match="each element of input_values must be a numpy array"):
            shm.set_shared_memory_region(h, [[1, 2]])
        assert shm.get_contents_as_numpy(h, np.float32, [0]).shape == (0,)
    finally:
        shm.destroy_shared_memory_region(h)
    assert shm.mapped_shared_memory_regions() == []
    assert str(shm.SharedMemoryException(__import__("ctypes").c_int(-5))) == "unable to unlink the shared memory region"
    with pytest.raises(shm.SharedMemoryException, match="unable to get shared memory descriptor"):
        shm.create_shared_memory_region("bad", "/no/such/dir/key", 8)


def test_bytes_contents_host_loop_and_quirks():
    """Small BYTES regions take the reference's loop: exactly prod(shape) elements, never fewer than one."""
    elems = [b"0", b"11", b"", b"\x00\x01\x02"]
    wire = b"".join(struct.pack("<I", len(x)) + x for x in elems)
    h = shm.create_shared_memory_region("strs", _key("bytes"), len(wire) + 8)
    try:
        shm.set_shared_memory_region(h, [np.asarray(wire, dtype=np.object_)])  # the 0-d serialized array
        got = shm.get_contents_as_numpy(h, np.object_, [2, 2])
        assert got.dtype == np.object_ and got.shape == (2, 2) and list(got.ravel()) == elems
        assert list(shm.get_contents_as_numpy(h, np.object_, [2])) == elems[:2]
        assert shm.get_contents_as_numpy(h, np.object_, []).item() == b"0"  # shape [] reads one element
        assert list(shm.get_contents_as_numpy(h, np.object_, [1], offset=5)) == [b"11"]
        with pytest.raises(struct.error):
            shm.get_contents_as_numpy(h, np.object_, [50])
    finally:
        shm.destroy_shared_memory_region(h)


@pytest.mark.gpu
def test_device_bridge_and_gpu_bytes_decode():
    import torch

    from triton_client_b200 import utils as u

    blob, lens = cases.gbytes(1000)
    arr = cases.to_object_array(blob, lens)
    ser = u.serialize_byte_tensor(arr)  # 0-d object array, as the examples pass it
    nbytes = len(ser.item())
    h = shm.create_shared_memory_region("gpu_strs", _key("gpu"), nbytes + 4096)
    try:
        shm.set_shared_memory_region(h, [ser])
        back = shm.get_contents_as_numpy(h, np.object_, [40, 25])  # decoded by the CUDA kernels
        assert back.shape == (40, 25) and np.array_equal(back.ravel(), arr)
        with pytest.raises(struct.error):
            shm.get_contents_as_numpy(h, np.object_, [5000])
        # device -> region -> device
        t = torch.randn(300, device="cuda")
        n = shm.set_shared_memory_region_from_device(h, [t, t[:10].contiguous()], offset=nbytes)
        assert n == 1240
        assert np.array_equal(shm.get_contents_as_numpy(h, np.float32, [300], offset=nbytes), t.cpu().numpy())
        out = torch.empty(10, device="cuda")
        shm.get_contents_to_device(h, out, offset=nbytes + 1200)
        assert torch.equal(out, t[:10])
        with pytest.raises(shm.SharedMemoryException, match="exceeds region"):
            shm.set_shared_memory_region_from_device(h, [t], offset=nbytes + 4000)
        # a serialised tensor produced on the device lands in the system region without host objects
        d_ser, total = u.serialize_to_device(torch.from_numpy(blob).cuda(), torch.from_numpy(lens).cuda())
        shm.set_shared_memory_region_from_device(h, [d_ser[:total]])
        assert bytes(shm.get_contents_as_numpy(h, np.uint8, [total])) == ser.item()
    finally:
        shm.destroy_shared_memory_region(h)
