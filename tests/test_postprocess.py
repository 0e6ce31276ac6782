"""Classification postprocess (the step after the BYTES decode):
the oracle against the goldens made by the reference example's own postprocess()
(/root/reference/src/python/examples/image_client.py:196-216, stdout captured by
tests/golden/make_golden.py), and the drop-in - decode + field split on the device - against both.
"""
import contextlib
import io
import json
import os
import struct

import numpy as np
import pytest

from oracle import postprocess as opp
from tests.golden import cases

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = {c["name"]: c for c in json.load(open(os.path.join(HERE, "golden", "postprocess_golden.json")))["cases"]}
CASES = cases.postprocess_cases()


def _run(fn):
    """(lines printed, [exception type name, message] or None)"""
    buf, err = io.StringIO(), None
    try:
        with contextlib.redirect_stdout(buf):
            fn()
    except Exception as e:
        err = [type(e).__name__, str(e)]
    text = buf.getvalue()
    return (text.split("\n")[:-1] if text else []), err


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_matches_reference_golden(name):
    a, batch, batching = CASES[name]

    def go():
        for line in opp.lines(a, batch, batching):
            print(line)

    lines, err = _run(go)
    if GOLD[name]["raises"]:
        assert err == GOLD[name]["raises"]
        # the oracle returns a list: nothing is printed before it raises; the lines before the bad
        # element are checked on the drop-in below
    else:
        assert err is None and lines == GOLD[name]["lines"]


def test_oracle_matches_reference_live():
    from oracle import reference_loader as rl

    if not rl.available():
        pytest.skip("/root/reference not present")
    post = rl.reference_postprocess()

    class R:
        def __init__(self, a):
            self.a = a

        def as_numpy(self, name):
            return self.a

    for name, (a, batch, batching) in CASES.items():
        lines, err = _run(lambda: post(R(a), "OUT", batch, batching))
        assert lines == GOLD[name]["lines"] and err == GOLD[name]["raises"], name


def _response(a):
    """An HTTP inference response carrying `a` as a binary BYTES output named OUT."""
    from triton_client_b200.http import InferenceServerClient

    sec = b"".join(struct.pack("<I", len(x)) + x for x in a.ravel().tolist())
    hdr = json.dumps({"model_name": "m", "model_version": "1", "outputs": [
        {"name": "OUT", "datatype": "BYTES", "shape": list(a.shape), "parameters": {"binary_data_size": len(sec)}}]},
        separators=(",", ":")).encode()
    return InferenceServerClient.parse_response_body(hdr + sec, header_length=len(hdr))


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_dropin_prints_what_the_reference_prints(name):
    from triton_client_b200 import _lib
    from triton_client_b200.image_client import postprocess

    a, batch, batching = CASES[name]
    res = _response(a)
    before = _lib.kernel_launch_count()
    lines, err = _run(lambda: postprocess(res, "OUT", batch, batching))
    assert lines == GOLD[name]["lines"]
    assert err == GOLD[name]["raises"]
    if name != "batch_mismatch_raises":
        assert _lib.kernel_launch_count() > before  # decode + field split ran on the device


@pytest.mark.gpu
def test_fields_of_a_large_tensor_equal_python_split():
    import torch

    from oracle import clib
    from triton_client_b200 import utils as u
    from triton_client_b200.image_client import classification_fields

    rng = np.random.default_rng(5)
    n = 200_000
    lens = rng.integers(0, 80, n)
    blob = rng.integers(0, 256, int(lens.sum()), dtype=np.uint8)
    blob[rng.random(blob.size) < 0.08] = ord(":")
    starts = np.concatenate(([0], np.cumsum(lens)))
    ser = clib.bytes_serialize(blob, lens.astype(np.int64))
    offsets, payload = u.deserialize_from_device(torch.from_numpy(np.asarray(ser)).cuda())
    got = classification_fields(offsets, payload).cpu().numpy()
    raw = blob.tobytes()
    want = np.empty((n, 4), dtype=np.int64)
    for i in range(n):
        e = raw[starts[i]:starts[i + 1]]
        pos, p = [], -1
        for _ in range(3):
            p = e.find(b":", p + 1)
            if p < 0:
                break
            pos.append(p)
        want[i] = pos + [-1] * (3 - len(pos)) + [len(e)]
    assert np.array_equal(got, want)


def test_host_path_for_non_bytes_outputs_matches_oracle():
    """Outputs that are not binary BYTES (here: a numpy str array) take the reference's host route."""
    from triton_client_b200.image_client import postprocess_lines

    class R:
        def as_numpy(self, name):
            return np.array([["0.5:1:a", "0.25:2:b:c"]])

    assert postprocess_lines(R(), "OUT", 1, True) == opp.lines(R().as_numpy("OUT"), 1, True)


@pytest.mark.gpu
def test_large_output_takes_the_device_route_and_prints_the_same():
    """More than 256 KB of classification strings: decode and split stay on the device
    (InferResult.as_device); same lines as the oracle."""
    from triton_client_b200.image_client import postprocess_lines
    from triton_client_b200.utils import _small

    rng = np.random.default_rng(9)
    rows, topk = 4000, 5
    a = np.empty((rows, topk), dtype=np.object_)
    for i in range(rows):
        for j in range(topk):
            a[i, j] = ("%.6f:%d:class %d of row %d" % (rng.random(), rng.integers(0, 1000), j, i)).encode()
    res = _response(a)
    assert len(res._binary_section(res.get_output("OUT"))[0]) > _small.MAX_BYTES
    assert postprocess_lines(res, "OUT", rows, True) == opp.lines(a, rows, True)
