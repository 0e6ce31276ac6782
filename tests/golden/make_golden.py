#!/usr/bin/env python
"""Generates tests/golden/*.json|*.npz by running the UNMODIFIED reference imported from
/root/reference (only possible in the build container; the GPU box has no /root/reference).

    python tests/golden/make_golden.py

Inputs are seeded numpy generators so that every test can rebuild them without the reference;
outputs are stored as sha256 hex digests (large cases) or as full arrays (tiny cases).
Recorded environment: Pillow / numpy versions of the generating container (Pillow's resampler is
third-party and unpinned by the reference; see oracle/__init__.py).
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import reference_loader as rl  # noqa: E402
from tests.golden import cases  # noqa: E402


def sha(b):
    return hashlib.sha256(b).hexdigest()


def main():
    import PIL
    from PIL import Image

    assert rl.available(), "needs /root/reference"
    pre = rl.reference_preprocess()
    utils = rl.reference_utils()
    env = {"pillow": PIL.__version__, "numpy": np.__version__}

    # ---- preprocess -------------------------------------------------------------------------
    gold = {"env": env, "cases": []}
    small = {}
    for case in cases.PREPROCESS_CASES:
        src = cases.make_image(case)
        img = cases.to_pil(src, case)
        out = pre(img, case["format"], case["dtype"], case["c"], case["h"], case["w"],
                  case["scaling"], "grpc")
        outc = np.ascontiguousarray(out)
        rec = dict(case)
        rec["src_sha256"] = sha(np.ascontiguousarray(src).tobytes())
        rec["out_sha256"] = sha(outc.tobytes())
        rec["out_dtype"] = str(outc.dtype)
        rec["out_shape"] = list(outc.shape)
        conv = img.convert("L" if case["c"] == 1 else "RGB")
        u8 = np.array(conv.resize((case["w"], case["h"]), Image.BILINEAR))
        rec["u8_sha256"] = sha(np.ascontiguousarray(u8).tobytes())
        gold["cases"].append(rec)
        if outc.nbytes <= 16384:
            small[case["name"]] = outc
    with open(os.path.join(HERE, "preprocess_golden.json"), "w") as f:
        json.dump(gold, f, indent=1)
    # ---- source modes that Pillow converts before the resize (CMYK, LA, P, YCbCr) --------------
    mg = {"env": env, "cases": []}
    for case in cases.MODE_CASES:
        src = cases.make_image(case)
        img = cases.to_pil(src, case)
        outc = np.ascontiguousarray(pre(img, case["format"], case["dtype"], case["c"], case["h"], case["w"],
                                        case["scaling"], "grpc"))
        conv = np.asarray(img.convert("L" if case["c"] == 1 else "RGB"))
        rec = dict(case)
        rec.update({"src_sha256": sha(np.ascontiguousarray(src).tobytes()), "converted_sha256": sha(np.ascontiguousarray(conv).tobytes()),
                    "out_sha256": sha(outc.tobytes()), "out_dtype": str(outc.dtype), "out_shape": list(outc.shape)})
        mg["cases"].append(rec)
    with open(os.path.join(HERE, "preprocess_modes_golden.json"), "w") as f:
        json.dump(mg, f, indent=1)
    np.savez_compressed(os.path.join(HERE, "preprocess_small.npz"), **small)

    # ---- BYTES ------------------------------------------------------------------------------
    bg = {"env": env, "gbytes": [], "objects": []}
    for n in cases.GBYTES_N:
        blob, lens = cases.gbytes(n)
        arr = cases.to_object_array(blob, lens)
        ser = utils.serialize_byte_tensor(arr).item()
        back = utils.deserialize_bytes_tensor(ser)
        assert np.array_equal(back, arr)
        bg["gbytes"].append({"n": n, "seed": 42, "payload_bytes": int(lens.sum()),
                             "serialized_bytes": len(ser), "blob_sha256": sha(blob.tobytes()),
                             "serialized_sha256": sha(ser)})
    # odd object/bytes_ inputs: what the reference's per-element normalisation yields
    for name, arr in cases.object_cases().items():
        r = utils.serialize_byte_tensor(arr)
        ser = r.item() if r.size > 0 else b""
        bg["objects"].append({"name": name, "serialized_hex": ser.hex(),
                              "result_shape": list(r.shape), "result_dtype": str(r.dtype)})
    with open(os.path.join(HERE, "bytes_golden.json"), "w") as f:
        json.dump(bg, f, indent=1)
    # ---- BF16 -------------------------------------------------------------------------------
    bf = {"env": env, "cases": []}
    for name, arr in cases.bf16_cases().items():
        r = utils.serialize_bf16_tensor(arr)
        ser = r.item()
        back = utils.deserialize_bf16_tensor(ser)
        bf["cases"].append({"name": name, "n": int(arr.size), "serialized_sha256": sha(ser),
                            "serialized_hex": ser.hex() if len(ser) <= 64 else None,
                            "result_shape": list(r.shape), "result_dtype": str(r.dtype),
                            "back_shape": list(back.shape), "back_dtype": str(back.dtype),
                            "back_sha256": sha(np.ascontiguousarray(back).tobytes())})
    with open(os.path.join(HERE, "bf16_golden.json"), "w") as f:
        json.dump(bf, f, indent=1)
    # ---- HTTP request bodies / response parsing (reference http module behind transport stand-ins) -----
    import gzip

    h = rl.reference_http()
    hg = {"env": env, "note": "rapidjson is absent from the image: the reference ran with a stand-in that writes "
                              "JSON compactly like rapidjson's default", "requests": [], "responses": []}
    for name, thunk in cases.http_request_cases(h).items():
        inputs, outputs, kw = thunk()
        body, json_size = h.InferenceServerClient.generate_request_body(inputs, outputs=outputs, **kw)
        hg["requests"].append({"name": name, "json_size": json_size, "body_sha256": sha(body), "body_len": len(body),
                               "header": body[:json_size].decode() if json_size else body.decode()})
    body, hl = cases.http_response_body()
    for enc in (None, "gzip"):
        payload = gzip.compress(body) if enc == "gzip" else body
        res = h.InferenceServerClient.parse_response_body(payload, header_length=hl, content_encoding=enc)
        rec = {"encoding": enc, "outputs": []}
        for out in res.get_response()["outputs"]:
            arr = res.as_numpy(out["name"])
            rec["outputs"].append({"name": out["name"], "dtype": str(arr.dtype), "shape": list(arr.shape),
                                   "items": [x.hex() if isinstance(x, bytes) else repr(x) for x in arr.ravel().tolist()]})
        hg["responses"].append(rec)
    with open(os.path.join(HERE, "http_golden.json"), "w") as f:
        json.dump(hg, f, indent=1)
    # ---- classification postprocess (the reference example's own function, stdout captured) -------------
    import contextlib
    import io

    post = rl.reference_postprocess()

    class _Result:
        def __init__(self, a):
            self.a = a

        def as_numpy(self, name):
            return self.a

    pg = {"env": env, "cases": []}
    for name, (a, batch, batching) in cases.postprocess_cases().items():
        buf, err = io.StringIO(), None
        try:
            with contextlib.redirect_stdout(buf):
                post(_Result(a), "OUT", batch, batching)
        except Exception as e:  # the reference raises for malformed elements / a wrong row count
            err = [type(e).__name__, str(e)]
        text = buf.getvalue()  # split on "\n" only: str.splitlines() also cuts at \x85 and friends
        pg["cases"].append({"name": name, "lines": text.split("\n")[:-1] if text else [], "raises": err})
    with open(os.path.join(HERE, "postprocess_golden.json"), "w") as f:
        json.dump(pg, f, indent=1)
    print("wrote goldens:", len(gold["cases"]), "preprocess cases,", len(bg["gbytes"]), "gbytes,",
          len(bg["objects"]), "object cases")


if __name__ == "__main__":
    main()
