"""triton_client_b200: the Triton client's tensor-marshalling path on B200 (sm_100a).

Drop-in modules (same names and signatures as the reference's):

* ``triton_client_b200.utils``                      <- ``tritonclient.utils`` (BYTES codec, dtype maps)
* ``triton_client_b200.utils.cuda_shared_memory``   <- ``tritonclient.utils.cuda_shared_memory``
* ``triton_client_b200.image_client.preprocess``    <- ``image_client.py::preprocess``

Everything computes in hand-written CUDA behind the C ABI of ``libtcb200.so``
(``include/tcb200.h``). Importing this package fails loudly if that library has not been built.
"""
from . import _lib  # noqa: F401  (raises ImportError when libtcb200.so is missing)

__version__ = "0.1.0"
