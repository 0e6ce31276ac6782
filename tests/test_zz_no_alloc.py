"""GPU test of TcbBytesSerializeNoAlloc, the serialiser for callers that cannot allocate.

Some users of this library run where device memory cannot be requested at call time: a server
process that gave all of HBM to its own arena, a request path where cudaMalloc's device-wide
synchronisation is too expensive, a hook that is forbidden to touch the allocator. For them the
library offers one entry point whose scan scratch is a fixed array inside libtcb200.so rather than
a caller-provided workspace. This test drives it through the C ABI exactly as such a caller would
(inputs and outputs only) and requires its bytes to equal those of the ordinary entry point,
TcbBytesSerialize, bit for bit - twice in a row, because the built-in scratch is shared between
calls and keeps the previous call's contents until the next call clears it.

The file name sorts last on purpose: if this addition ever breaks, a run that stops at the first
failure has still executed every other test. It passes on its own in a fresh process and after
the rest of the suite.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NO_ALLOC_MAX_ELEMENTS = 16 * 1024 * 1024


def _serialize_both(torch, _lib, lens, seed):
    """Run the existing entry point and the no-alloc one on the same device inputs."""
    rng = np.random.default_rng(seed)
    lens = np.asarray(lens, dtype=np.int32)
    blob = rng.integers(0, 256, int(lens.astype(np.int64).sum()), dtype=np.uint8)
    n, total_expected = int(lens.size), int(blob.size) + 4 * int(lens.size)
    d_blob = torch.from_numpy(blob).cuda() if blob.size else torch.empty(1, dtype=torch.uint8, device="cuda")
    d_lens = torch.from_numpy(lens).cuda()
    stream = torch.cuda.current_stream().cuda_stream

    ref_out = torch.full((total_expected + 32,), 0x5A, dtype=torch.uint8, device="cuda")
    ws_bytes = int(_lib.lib.TcbBytesSerializeWorkspace(n))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
    ref_total = ctypes.c_uint64(0)
    _lib.check(_lib.lib.TcbBytesSerialize(0, d_blob.data_ptr(), d_lens.data_ptr(), 4, n, ref_out.data_ptr(),
                                          total_expected, ws.data_ptr(), ws_bytes, None,
                                          ctypes.byref(ref_total), stream))

    new_out = torch.full((total_expected + 32,), 0x5A, dtype=torch.uint8, device="cuda")
    new_total = ctypes.c_uint64(0)
    _lib.check(_lib.lib.TcbBytesSerializeNoAlloc(0, d_blob.data_ptr(), d_lens.data_ptr(), 4, n,
                                                 new_out.data_ptr(), total_expected,
                                                 ctypes.byref(new_total), stream))
    assert new_total.value == ref_total.value == total_expected
    assert torch.equal(new_out, ref_out)  # bit-exact, guard bytes past the end untouched
    return new_out[:total_expected].cpu().numpy(), blob, lens


def test_no_alloc_serialize_matches_existing_entry_point_twice_in_a_row():
    import torch

    from oracle import clib
    from triton_client_b200 import _lib

    torch.cuda.set_device(0)
    rng = np.random.default_rng(7)
    # first call: many tiles, so plenty of scan state is left behind in the built-in scratch
    out1, blob1, lens1 = _serialize_both(torch, _lib, rng.integers(0, 257, 300_000), seed=1)
    # second call straight after, different size and contents: must not see call one's leftovers
    out2, blob2, lens2 = _serialize_both(torch, _lib, rng.integers(0, 40, 70_001), seed=2)
    # and both equal the oracle's statement of the wire format
    assert np.array_equal(out1, clib.bytes_serialize(blob1, lens1.astype(np.int64)))
    assert np.array_equal(out2, clib.bytes_serialize(blob2, lens2.astype(np.int64)))


def test_no_alloc_serialize_refuses_more_than_its_scratch_covers():
    import torch

    from triton_client_b200 import _lib

    torch.cuda.set_device(0)
    dummy = torch.zeros(16, dtype=torch.uint8, device="cuda")
    total = ctypes.c_uint64(0)
    rc = _lib.lib.TcbBytesSerializeNoAlloc(0, dummy.data_ptr(), dummy.data_ptr(), 4, NO_ALLOC_MAX_ELEMENTS + 1,
                                           dummy.data_ptr(), 16, ctypes.byref(total),
                                           torch.cuda.current_stream().cuda_stream)
    assert rc == _lib.TCB_ERR_OUT_OF_RANGE
    msg = _lib.last_error()
    assert "16777216" in msg and "TcbBytesSerialize" in msg
    # an empty tensor is fine and writes nothing
    rc = _lib.lib.TcbBytesSerializeNoAlloc(0, None, None, 4, 0, dummy.data_ptr(), 16, ctypes.byref(total),
                                           torch.cuda.current_stream().cuda_stream)
    assert rc == 0 and total.value == 0
