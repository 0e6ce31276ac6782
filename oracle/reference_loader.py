"""Oracle (test infrastructure): import the UNMODIFIED reference from /root/reference.

Only usable in the build container (the GPU box has no /root/reference); used by
tests/golden/make_golden.py to generate fixtures and by CPU tests (skipped when absent) to
validate the restatements. Method from SURVEY.md 8c: `tritonclient.utils` needs only numpy;
`image_client.preprocess` imports after stubbing the three transport modules it pulls in.
"""
import importlib.util
import os
import sys
import types

REFERENCE_ROOT = "/root/reference"
_LIB = os.path.join(REFERENCE_ROOT, "src/python/library")
_IMAGE_CLIENT = os.path.join(REFERENCE_ROOT, "src/python/examples/image_client.py")


def available():
    return os.path.isdir(_LIB) and os.path.isfile(_IMAGE_CLIENT)


def reference_utils():
    # appended, not prepended: the reference tree has its own top-level `tests` package, which must
    # not shadow this repo's (spawned gloo ranks re-import tests.* with the parent's sys.path)
    if _LIB not in sys.path:
        sys.path.append(_LIB)
    import tritonclient.utils as u  # noqa
    return u


def reference_preprocess():
    reference_utils()
    for name in ("tritonclient.grpc", "tritonclient.http", "tritonclient.grpc.model_config_pb2"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    mc = sys.modules["tritonclient.grpc.model_config_pb2"]
    if not hasattr(mc, "ModelInput"):
        mi = types.SimpleNamespace(FORMAT_NONE=0, FORMAT_NHWC=1, FORMAT_NCHW=2)
        mc.ModelInput = mi
    sys.modules["tritonclient.grpc"].model_config_pb2 = mc
    return _image_client_module().preprocess


def _image_client_module():
    spec = importlib.util.spec_from_file_location("_ref_image_client", _IMAGE_CLIENT)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def reference_postprocess():
    """The reference example's postprocess(results, output_name, batch_size, supports_batching)."""
    reference_preprocess()  # installs the transport stand-ins
    return _image_client_module().postprocess


def reference_http():
    """The reference's HTTP client module (InferInput, InferRequestedOutput, InferResult,
    InferenceServerClient.generate_request_body / parse_response_body). Its transport imports are
    absent from this image, so stand-ins are installed first: ``rapidjson`` -> the standard json
    module writing compactly (rapidjson's default output has no whitespace), ``gevent`` /
    ``geventhttpclient`` -> empty modules (never called by the two server-free static methods)."""
    import json as _json

    reference_utils()
    if "rapidjson" not in sys.modules:
        rj = types.ModuleType("rapidjson")
        rj.dumps = lambda obj, **kw: _json.dumps(obj, separators=(",", ":"))
        rj.loads = _json.loads
        sys.modules["rapidjson"] = rj
    for name in ("gevent", "gevent.pool", "geventhttpclient", "geventhttpclient.url"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["gevent"].pool = sys.modules["gevent.pool"]
    sys.modules["geventhttpclient"].HTTPClient = object
    sys.modules["geventhttpclient.url"].URL = object
    sys.modules["geventhttpclient"].url = sys.modules["geventhttpclient.url"]
    for name in ("tritonclient.http", "tritonclient.grpc"):  # drop the stubs reference_preprocess() may have left
        m = sys.modules.get(name)
        if m is not None and not hasattr(m, "__path__") and not hasattr(m, "InferInput"):
            del sys.modules[name]
    import tritonclient.http as h  # noqa
    return h
