"""CPU oracle for the tensor-marshalling hot path. TEST INFRASTRUCTURE ONLY.

Nothing in the product package (``triton_client_b200``) imports this package.
Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` leg may import it, and there only as the checker or the timed
CPU baseline, never as the thing shipped.

What it restates (every function cites the reference file:line it follows):

* ``oracle.pil_bilinear`` - Pillow's 8-bit BILINEAR resampler (third-party, Pillow
  12.2.0 ``src/libImaging/Resample.c``; Pillow is *not* under ``/root/reference`` and is
  *unpinned* by the reference) and the reference's
  ``image_client.py::preprocess`` (``src/python/examples/image_client.py:154-193``).
* ``oracle.bytes_codec`` - ``tritonclient.utils.serialize_byte_tensor`` /
  ``deserialize_bytes_tensor`` / ``serialized_byte_size``
  (``src/python/library/tritonclient/utils/__init__.py:43-68,193-276``).
* ``oracle/csrc/oracle.c`` - the same two algorithms in plain C (built by
  ``oracle/Makefile`` into ``oracle/liboracle.so``); used for larger parity cases and as
  the threaded CPU port when timing.

Parity pinning:

* BYTES path: PINNED. Checked against the reference's own known-answer vectors
  (``src/c++/tests/cc_client_test.cc:1929-1953,2140-2153``;
  ``src/python/library/tests/test_cuda_shared_memory.py:137-164``;
  ``src/python/examples/simple_grpc_string_infer_client.py:168-180``) and against outputs
  of the reference itself imported from ``/root/reference`` (fixtures in
  ``tests/golden/``, generator ``tests/golden/make_golden.py``).
* preprocess path: the reference holds NO test or golden vector for ``preprocess`` and
  does not pin Pillow, so against the reference's own tests this parity is UNPINNED. It
  is pinned here against outputs of the reference's ``preprocess`` itself, imported
  unmodified from ``/root/reference`` and run in the build container with Pillow 12.2.0
  / numpy 2.3.5 (``tests/golden/preprocess_*.npz``; sha256 prefixes from SURVEY.md 8c).
"""
