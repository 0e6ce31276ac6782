"""The callers of the marshalling path: ``InferInput``, ``InferRequestedOutput``,
``InferResult`` and the two server-free entry points of the HTTP client,
``InferenceServerClient.generate_request_body`` / ``parse_response_body``.

Same names, argument meaning, defaults, error texts and wire bytes as
/root/reference/src/python/library/tritonclient/http/_infer_input.py:36-272,
_requested_output.py:36-117, _infer_result.py:36-242, _utils.py:77-150 and
_client.py:1218-1329 (KServe v2 JSON header + binary tensor data). The transport (connection pool,
gevent, TLS, the server calls) is out of scope and not here.

What is new is where the tensor bytes come from: BYTES and BF16 tensors are (de)serialised by the
CUDA kernels of libtcb200.so, and tensors that already live on the device never take the
reference's detour through per-element Python objects:

* ``InferInput.set_data_from_device`` - device payload + lengths (BYTES) / device float32 (BF16) /
  any device tensor: one kernel + one D2H copy produce the request's binary section;
* ``InferInput.set_data_into_shared_memory`` - the tensor is serialised or copied straight into a
  registered CUDA region at a byte offset and the input refers to it (``shared_memory_offset``, so
  many tensors or a whole batch share one region); nothing serialized ever exists on the host;
* ``InferResult.as_numpy_from_shared_memory`` / ``as_device`` - outputs decoded from a region or
  into device tensors.

The JSON header is written compactly (no whitespace), as the reference's rapidjson does.
"""
import gzip
import json
import zlib

import numpy as np

from .._device import to_host
from ..utils import (InferenceServerException, deserialize_bf16_tensor, deserialize_bytes_tensor,
                     np_to_triton_dtype, raise_error, serialize_bf16_tensor, serialize_byte_tensor,
                     triton_to_np_dtype)
from .. import utils as _utils

__all__ = ["InferenceServerClient", "InferInput", "InferRequestedOutput", "InferResult",
           "InferenceServerException"]

_FIXED_ITEMSIZE = {"BOOL": 1, "INT8": 1, "INT16": 2, "INT32": 4, "INT64": 8, "UINT8": 1, "UINT16": 2,
                   "UINT32": 4, "UINT64": 8, "FP16": 2, "FP32": 4, "FP64": 8}


def _dumps(obj):
    return json.dumps(obj, separators=(",", ":"))


class InferInput:
    """Describes one input tensor of an inference request (http/_infer_input.py:36-272)."""

    def __init__(self, name, shape, datatype):
        self._name = name
        self._shape = shape
        self._datatype = datatype
        self._parameters = {}
        self._data = None
        self._raw_data = None

    def name(self):
        return self._name

    def datatype(self):
        return self._datatype

    def shape(self):
        return self._shape

    def set_shape(self, shape):
        self._shape = shape
        return self

    def _check_shape(self, got_shape):
        valid_shape = len(self._shape) == len(got_shape) and all(
            int(self._shape[i]) == int(got_shape[i]) for i in range(len(self._shape)))
        if not valid_shape:
            raise_error("got unexpected numpy array shape [{}], expected [{}]".format(
                str(tuple(got_shape))[1:-1], str(self._shape)[1:-1]))

    def _clear_shared_memory(self):
        self._parameters.pop("shared_memory_region", None)
        self._parameters.pop("shared_memory_byte_size", None)
        self._parameters.pop("shared_memory_offset", None)

    def set_data_from_numpy(self, input_tensor, binary_data=True):
        """Same contract as the reference (:106-242); BYTES / BF16 tensors are serialised on the GPU."""
        if not isinstance(input_tensor, (np.ndarray,)):
            raise_error("input_tensor must be a numpy array")
        if self._datatype == "BF16":
            if input_tensor.dtype != triton_to_np_dtype(self._datatype):
                raise_error("got unexpected datatype {} from numpy array, expected {} for BF16 type".format(
                    input_tensor.dtype, triton_to_np_dtype(self._datatype)))
        else:
            dtype = np_to_triton_dtype(input_tensor.dtype)
            if self._datatype != dtype:
                raise_error("got unexpected datatype {} from numpy array, expected {}".format(dtype, self._datatype))
        self._check_shape(input_tensor.shape)
        self._clear_shared_memory()
        if not binary_data:
            self._parameters.pop("binary_data_size", None)
            self._raw_data = None
            if self._datatype == "BF16":
                raise_error("BF16 inputs must be sent as binary data over HTTP. Please set binary_data=True")
            if self._datatype == "BYTES":
                self._data = []
                try:
                    if input_tensor.size > 0:
                        for obj in np.nditer(input_tensor, flags=["refs_ok"], order="C"):
                            if input_tensor.dtype == np.object_:
                                if type(obj.item()) == bytes:
                                    self._data.append(str(obj.item(), encoding="utf-8"))
                                else:
                                    self._data.append(str(obj.item()))
                            else:
                                self._data.append(str(obj.item(), encoding="utf-8"))
                except UnicodeDecodeError:
                    raise_error(f'Failed to encode "{obj.item()}" using UTF-8. Please use binary_data=True, if'
                                " you want to pass a byte array.")
            else:
                self._data = [val.item() for val in input_tensor.flatten()]
        else:
            self._data = None
            if self._datatype == "BYTES":
                serialized_output = serialize_byte_tensor(input_tensor)
                self._raw_data = serialized_output.item() if serialized_output.size > 0 else b""
            elif self._datatype == "BF16":
                serialized_output = serialize_bf16_tensor(input_tensor)
                self._raw_data = serialized_output.item() if serialized_output.size > 0 else b""
            else:
                self._raw_data = input_tensor.tobytes()
            self._parameters["binary_data_size"] = len(self._raw_data)
        return self

    def set_shared_memory(self, region_name, byte_size, offset=0):
        """Refer to tensor data in a registered region (:244-270)."""
        self._data = None
        self._raw_data = None
        self._parameters.pop("binary_data_size", None)
        self._parameters["shared_memory_region"] = region_name
        self._parameters["shared_memory_byte_size"] = byte_size
        if offset != 0:
            self._parameters["shared_memory_offset"] = offset
        return self

    # ---- device-resident sources (no counterpart in the reference) ------------------------------
    def _device_wire(self, values, lengths):
        """(device uint8 tensor holding this tensor's wire bytes, byte count) for device inputs."""
        import torch

        if self._datatype == "BYTES":
            if lengths is None:
                raise_error("BYTES input from the device needs the payload and the per-element lengths")
            n = 1
            for s in self._shape:
                n *= int(s)
            if int(lengths.numel()) != n:
                raise_error("got {} lengths, expected {} elements for shape [{}]".format(
                    int(lengths.numel()), n, str(self._shape)[1:-1]))
            out, total = _utils.serialize_to_device(values, lengths)
            return out, int(total)
        if self._datatype == "BF16":
            self._check_shape(values.shape)
            out, total = _utils.serialize_bf16_to_device(values.contiguous())
            return out, int(total)
        np_dt = triton_to_np_dtype(self._datatype)
        if np_dt is None or self._datatype not in _FIXED_ITEMSIZE:
            raise_error("unsupported datatype {}".format(self._datatype))
        if values.element_size() != _FIXED_ITEMSIZE[self._datatype]:
            raise_error("got unexpected element size {} from the device tensor, expected {}".format(
                values.element_size(), self._datatype))
        self._check_shape(values.shape)
        flat = values.contiguous().view(torch.uint8).reshape(-1) if values.numel() else values.new_empty(0, dtype=torch.uint8)
        return flat, int(flat.numel())

    def set_data_from_device(self, values, lengths=None):
        """Binary tensor data from device memory. BYTES: ``values`` is the uint8 payload blob (all
        elements back to back, C order) and ``lengths`` the int32/int64 length of every element;
        BF16: a float32 CUDA tensor; other datatypes: a CUDA tensor of that type and shape. The wire
        bytes are produced on the device and read back with one copy."""
        wire, nbytes = self._device_wire(values, lengths)
        self._clear_shared_memory()
        self._data = None
        dev = wire.device.index
        self._raw_data = to_host(wire.data_ptr(), nbytes, dev).tobytes() if nbytes else b""
        self._parameters["binary_data_size"] = len(self._raw_data)
        return self

    def set_data_into_shared_memory(self, region, values, lengths=None, offset=0, region_name=None):
        """Write this tensor's wire bytes straight into a registered CUDA region at ``offset`` and
        refer to them (``shared_memory_region`` / ``_byte_size`` / ``_offset``). ``values`` may be a
        device tensor (see ``set_data_from_device``) or a numpy array. Returns the byte count, so the
        next tensor can be placed behind this one in the same region."""
        import torch

        from ..utils import cuda_shared_memory as cudashm

        dev = region._device_id
        if isinstance(values, np.ndarray):
            if self._datatype == "BYTES":
                nbytes = int(_utils.serialize_byte_tensor_to(region, values, offset))
            elif self._datatype == "BF16":
                t = torch.from_numpy(np.ascontiguousarray(values, dtype=np.float32)).to("cuda:%d" % dev)
                self._check_shape(values.shape)
                _, nbytes = _utils.serialize_bf16_to_device(t, region, offset, dev)
            else:
                self._check_shape(values.shape)
                cudashm.set_shared_memory_region(region, [values], offset=offset)
                nbytes = int(values.nbytes)
        elif self._datatype == "BYTES":
            _, nbytes = _utils.serialize_to_device(values, lengths, region, offset, dev)
        elif self._datatype == "BF16":
            self._check_shape(values.shape)
            _, nbytes = _utils.serialize_bf16_to_device(values.contiguous(), region, offset, dev)
        else:
            self._check_shape(values.shape)
            cudashm.set_shared_memory_region_from_dlpack(region, [values.contiguous()], offset=offset)
            nbytes = int(values.numel() * values.element_size())
        self.set_shared_memory(region_name if region_name is not None else region._triton_shm_name, int(nbytes), offset)
        return int(nbytes)

    def _get_binary_data(self):
        return self._raw_data

    def _get_tensor(self):
        tensor = {"name": self._name, "shape": self._shape, "datatype": self._datatype}
        if self._parameters:
            tensor["parameters"] = self._parameters
        if self._parameters.get("shared_memory_region") is None and self._raw_data is None:
            if self._data is not None:
                tensor["data"] = self._data
        return tensor


class InferRequestedOutput:
    """Describes a requested output (http/_requested_output.py:36-117)."""

    def __init__(self, name, binary_data=True, class_count=0):
        self._name = name
        self._parameters = {}
        if class_count != 0:
            self._parameters["classification"] = class_count
        self._binary = binary_data
        self._parameters["binary_data"] = binary_data

    def name(self):
        return self._name

    def set_shared_memory(self, region_name, byte_size, offset=0):
        if "classification" in self._parameters:
            raise_error("shared memory can't be set on classification output")
        if self._binary:
            self._parameters["binary_data"] = False
        self._parameters["shared_memory_region"] = region_name
        self._parameters["shared_memory_byte_size"] = byte_size
        if offset != 0:
            self._parameters["shared_memory_offset"] = offset

    def unset_shared_memory(self):
        self._parameters["binary_data"] = self._binary
        self._parameters.pop("shared_memory_region", None)
        self._parameters.pop("shared_memory_byte_size", None)
        self._parameters.pop("shared_memory_offset", None)

    def _get_tensor(self):
        tensor = {"name": self._name}
        if self._parameters:
            tensor["parameters"] = self._parameters
        return tensor


_RESERVED = ("sequence_id", "sequence_start", "sequence_end", "priority", "binary_data_output")


def _get_inference_request(inputs, request_id, outputs, sequence_id, sequence_start, sequence_end, priority,
                           timeout, custom_parameters):
    """http/_utils.py:77-150: JSON header, then the binary sections of the inputs in order."""
    infer_request = {}
    parameters = {}
    if request_id != "":
        infer_request["id"] = request_id
    if sequence_id != 0 and sequence_id != "":
        parameters["sequence_id"] = sequence_id
        parameters["sequence_start"] = sequence_start
        parameters["sequence_end"] = sequence_end
    if priority != 0:
        parameters["priority"] = priority
    if timeout is not None:
        parameters["timeout"] = timeout
    infer_request["inputs"] = [this_input._get_tensor() for this_input in inputs]
    if outputs:
        infer_request["outputs"] = [this_output._get_tensor() for this_output in outputs]
    else:
        parameters["binary_data_output"] = True
    if custom_parameters:
        for key, value in custom_parameters.items():
            if key in _RESERVED:
                raise_error(f'Parameter "{key}" is a reserved parameter and cannot be specified.')
            else:
                parameters[key] = value
    if parameters:
        infer_request["parameters"] = parameters
    request_json = _dumps(infer_request)
    json_size = len(request_json)
    request_body = [request_json.encode()]
    for input_tensor in inputs:
        raw_data = input_tensor._get_binary_data()
        if raw_data is not None:
            request_body.append(raw_data)
    if len(request_body) == 1:
        return request_body[0], None
    return b"".join(request_body), json_size


class InferResult:
    """Holds the response of an inference request (http/_infer_result.py:36-242). ``response`` is any
    object with ``get(header_name)`` and ``read(length=-1)`` (the reference passes a geventhttpclient
    response; ``from_response_body`` wraps plain bytes)."""

    def __init__(self, response, verbose):
        header_length = response.get("Inference-Header-Content-Length")

        class DecompressedResponse:
            def __init__(self, decompressed_data):
                self.decompressed_data_ = decompressed_data
                self.offset_ = 0

            def read(self, length=-1):
                if length == -1:
                    return self.decompressed_data_[self.offset_:]
                prev_offset = self.offset_
                self.offset_ += length
                return self.decompressed_data_[prev_offset:self.offset_]

        content_encoding = response.get("Content-Encoding")
        if content_encoding is not None:
            if content_encoding == "gzip":
                response = DecompressedResponse(gzip.decompress(response.read()))
            elif content_encoding == "deflate":
                response = DecompressedResponse(zlib.decompress(response.read()))
        if header_length is None:
            content = response.read()
            if verbose:
                print(content)
            try:
                self._result = json.loads(content)
            except UnicodeDecodeError as e:
                raise_error(f"Failed to encode using UTF-8. Please use binary_data=True, if"
                            f" you want to pass a byte array. UnicodeError: {e}")
        else:
            header_length = int(header_length)
            content = response.read(length=header_length)
            if verbose:
                print(content)
            self._result = json.loads(content)
            self._output_name_to_buffer_map = {}
            self._buffer = response.read()
            buffer_index = 0
            for output in self._result["outputs"]:
                parameters = output.get("parameters")
                if parameters is not None:
                    this_data_size = parameters.get("binary_data_size")
                    if this_data_size is not None:
                        self._output_name_to_buffer_map[output["name"]] = buffer_index
                        buffer_index = buffer_index + this_data_size

    @classmethod
    def from_response_body(cls, response_body, verbose=False, header_length=None, content_encoding=None):
        class Response:
            def __init__(self, response_body, header_length, content_encoding):
                self.response_body_ = response_body
                self.offset_ = 0
                self.parameters_ = {"Inference-Header-Content-Length": header_length,
                                    "Content-Encoding": content_encoding}

            def get(self, key):
                return self.parameters_.get(key)

            def read(self, length=-1):
                if length == -1:
                    return self.response_body_[self.offset_:]
                prev_offset = self.offset_
                self.offset_ += length
                return self.response_body_[prev_offset:self.offset_]

        return cls(Response(response_body, header_length, content_encoding), verbose)

    def _binary_section(self, output):
        """(bytes of the output's binary section, True) or (None, False) when it is JSON data."""
        parameters = output.get("parameters")
        if parameters is not None:
            this_data_size = parameters.get("binary_data_size")
            if this_data_size is not None:
                if this_data_size == 0:
                    return b"", True
                start_index = self._output_name_to_buffer_map[output["name"]]
                return self._buffer[start_index:start_index + this_data_size], True
        return None, False

    def as_numpy(self, name):
        """The named output as a numpy array, or None (:157-210). BYTES / BF16 sections are decoded on
        the GPU."""
        if self._result.get("outputs") is not None:
            for output in self._result["outputs"]:
                if output["name"] == name:
                    datatype = output["datatype"]
                    section, has_binary_data = self._binary_section(output)
                    if has_binary_data:
                        if len(section) != 0:
                            if datatype == "BYTES":
                                np_array = deserialize_bytes_tensor(section)
                            elif datatype == "BF16":
                                np_array = deserialize_bf16_tensor(section)
                            else:
                                np_array = np.frombuffer(section, dtype=triton_to_np_dtype(datatype))
                        else:
                            np_array = np.empty(0)
                    if not has_binary_data:
                        np_array = np.array(output["data"], dtype=triton_to_np_dtype(datatype))
                    np_array = np_array.reshape(output["shape"])
                    return np_array
        return None

    # ---- device-side readers (no counterpart in the reference) ----------------------------------
    def as_device(self, name, device=None):
        """The named binary output on the device, without per-element host work: BYTES ->
        (offsets int64 [n + 1], payload uint8) CUDA tensors; BF16 -> float32 CUDA tensor; other
        datatypes -> a uint8 CUDA tensor of the raw bytes. None if the output is absent or was sent
        as JSON."""
        from .._device import current_device, to_device_u8

        output = self.get_output(name)
        if output is None:
            return None
        section, has_binary_data = self._binary_section(output)
        if not has_binary_data:
            return None
        dev = current_device() if device is None else device
        d = to_device_u8(section, dev) if len(section) else None
        if output["datatype"] == "BYTES":
            import torch

            if d is None:
                return torch.zeros(1, dtype=torch.int64, device="cuda:%d" % dev), torch.empty(0, dtype=torch.uint8, device="cuda:%d" % dev)
            return _utils.deserialize_from_device(d, len(section), -1, 0, dev)
        if output["datatype"] == "BF16" and d is not None:
            return _utils.deserialize_bf16_from_device(d, len(section), 0, None, dev)
        return d

    def as_numpy_from_shared_memory(self, name, region, offset=None):
        """The named output read from the registered CUDA region the request asked it to be written to:
        datatype and shape come from the response, the byte offset from the response's
        ``shared_memory_offset`` unless given. BYTES are decoded on the device and only the payload
        crosses PCIe (cuda_shared_memory/__init__.py:242-325 copies the whole region and loops)."""
        from ..utils import cuda_shared_memory as cudashm

        output = self.get_output(name)
        if output is None:
            return None
        if offset is None:
            offset = int((output.get("parameters") or {}).get("shared_memory_offset", 0))
        return cudashm.get_contents_as_numpy(region, triton_to_np_dtype(output["datatype"]), output["shape"],
                                             offset=offset)

    def get_output(self, name):
        for output in self._result["outputs"]:
            if output["name"] == name:
                return output
        return None

    def get_response(self):
        return self._result


class InferenceServerClient:
    """Only the two server-free static methods of the reference's HTTP client (http/_client.py:1218-1329);
    connections and server calls are out of this library's scope."""

    @staticmethod
    def generate_request_body(inputs, outputs=None, request_id="", sequence_id=0, sequence_start=False,
                              sequence_end=False, priority=0, timeout=None, parameters=None):
        """(request body bytes, JSON header size or None) for the given inputs / outputs."""
        return _get_inference_request(inputs=inputs, request_id=request_id, outputs=outputs, sequence_id=sequence_id,
                                      sequence_start=sequence_start, sequence_end=sequence_end, priority=priority,
                                      timeout=timeout, custom_parameters=parameters)

    @staticmethod
    def parse_response_body(response_body, verbose=False, header_length=None, content_encoding=None):
        return InferResult.from_response_body(response_body, verbose, header_length, content_encoding)
