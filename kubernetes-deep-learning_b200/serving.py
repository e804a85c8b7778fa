"""Wire-compatible stand-in for the TF-Serving container of the reference.

The unmodified gateway (`model_server.py`) opens an insecure gRPC channel to ``$TF_SERVING_HOST`` (model_server.py:13-16) and calls
``tensorflow.serving.PredictionService/Predict`` with model ``clothing-model``, signature ``serving_default``, input ``input_8``
as a DT_FLOAT TensorProto carrying ``tensor_content`` (model_server.py:35-43), then reads ``outputs['dense_7'].float_val``
(model_server.py:46-49).  This module serves exactly that RPC on top of the CUDA engine, so pointing ``TF_SERVING_HOST`` at it
needs no change on the gateway side.

tensorflow / tensorflow-serving-api are not installable here, so the four protobuf messages involved are declared at run time with
the field numbers of the upstream .proto files (tensorflow/core/framework/{tensor,tensor_shape,types}.proto,
tensorflow_serving/apis/{model,predict}.proto); unknown fields a client may send are ignored by protobuf.
"""
import threading
from concurrent import futures
from typing import Callable, Dict, Optional

import numpy as np
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

MODEL_NAME = "clothing-model"          # tf-serving.dockerfile:4, model_server.py:40
SIGNATURE_NAME = "serving_default"     # model_server.py:41
INPUT_KEY = "input_8"                  # model_server.py:42
OUTPUT_KEY = "dense_7"                 # model_server.py:47
SERVICE = "tensorflow.serving.PredictionService"

DT_FLOAT, DT_DOUBLE, DT_INT32, DT_UINT8 = 1, 2, 3, 4

_F = descriptor_pb2.FieldDescriptorProto


def _field(msg, name, number, ftype, label=_F.LABEL_OPTIONAL, type_name=None, packed=None):
    f = msg.field.add()
    f.name, f.number, f.type, f.label = name, number, ftype, label
    if type_name:
        f.type_name = type_name
    if packed is not None:
        f.options.packed = packed
    return f


def _build_pool():
    pool = descriptor_pool.DescriptorPool()

    # google/protobuf/wrappers.proto (only Int64Value is needed, for ModelSpec.version)
    w = descriptor_pb2.FileDescriptorProto(name="kdl/wrappers.proto", package="google.protobuf", syntax="proto3")
    m = w.message_type.add(name="Int64Value")
    _field(m, "value", 1, _F.TYPE_INT64)
    pool.Add(w)

    # tensorflow/core/framework: types.proto (enum subset), tensor_shape.proto, tensor.proto
    t = descriptor_pb2.FileDescriptorProto(name="kdl/tensor.proto", package="tensorflow", syntax="proto3")
    e = t.enum_type.add(name="DataType")
    for n, v in (("DT_INVALID", 0), ("DT_FLOAT", 1), ("DT_DOUBLE", 2), ("DT_INT32", 3), ("DT_UINT8", 4), ("DT_INT16", 5),
                 ("DT_INT8", 6), ("DT_STRING", 7), ("DT_INT64", 9), ("DT_BOOL", 10), ("DT_BFLOAT16", 14), ("DT_HALF", 19)):
        e.value.add(name=n, number=v)
    shp = t.message_type.add(name="TensorShapeProto")
    dim = shp.nested_type.add(name="Dim")
    _field(dim, "size", 1, _F.TYPE_INT64)
    _field(dim, "name", 2, _F.TYPE_STRING)
    _field(shp, "dim", 2, _F.TYPE_MESSAGE, _F.LABEL_REPEATED, ".tensorflow.TensorShapeProto.Dim")
    _field(shp, "unknown_rank", 3, _F.TYPE_BOOL)
    ten = t.message_type.add(name="TensorProto")
    _field(ten, "dtype", 1, _F.TYPE_ENUM, type_name=".tensorflow.DataType")
    _field(ten, "tensor_shape", 2, _F.TYPE_MESSAGE, type_name=".tensorflow.TensorShapeProto")
    _field(ten, "version_number", 3, _F.TYPE_INT32)
    _field(ten, "tensor_content", 4, _F.TYPE_BYTES)
    _field(ten, "float_val", 5, _F.TYPE_FLOAT, _F.LABEL_REPEATED, packed=True)
    _field(ten, "double_val", 6, _F.TYPE_DOUBLE, _F.LABEL_REPEATED, packed=True)
    _field(ten, "int_val", 7, _F.TYPE_INT32, _F.LABEL_REPEATED, packed=True)
    pool.Add(t)

    # tensorflow_serving/apis: model.proto, predict.proto
    s = descriptor_pb2.FileDescriptorProto(name="kdl/predict.proto", package="tensorflow.serving", syntax="proto3")
    s.dependency.extend(["kdl/wrappers.proto", "kdl/tensor.proto"])
    ms = s.message_type.add(name="ModelSpec")
    _field(ms, "name", 1, _F.TYPE_STRING)
    _field(ms, "version", 2, _F.TYPE_MESSAGE, type_name=".google.protobuf.Int64Value")
    _field(ms, "signature_name", 3, _F.TYPE_STRING)
    _field(ms, "version_label", 4, _F.TYPE_STRING)

    def add_map(msg, field_name, number):
        entry = msg.nested_type.add(name="".join(p.capitalize() for p in field_name.split("_")) + "Entry")
        entry.options.map_entry = True
        _field(entry, "key", 1, _F.TYPE_STRING)
        _field(entry, "value", 2, _F.TYPE_MESSAGE, type_name=".tensorflow.TensorProto")
        _field(msg, field_name, number, _F.TYPE_MESSAGE, _F.LABEL_REPEATED, f".tensorflow.serving.{msg.name}.{entry.name}")

    req = s.message_type.add(name="PredictRequest")
    _field(req, "model_spec", 1, _F.TYPE_MESSAGE, type_name=".tensorflow.serving.ModelSpec")
    add_map(req, "inputs", 2)
    _field(req, "output_filter", 3, _F.TYPE_STRING, _F.LABEL_REPEATED)
    res = s.message_type.add(name="PredictResponse")
    add_map(res, "outputs", 1)
    _field(res, "model_spec", 2, _F.TYPE_MESSAGE, type_name=".tensorflow.serving.ModelSpec")
    pool.Add(s)
    return pool


_POOL = _build_pool()


def _cls(name):
    return message_factory.GetMessageClass(_POOL.FindMessageTypeByName(name))


TensorProto = _cls("tensorflow.TensorProto")
PredictRequest = _cls("tensorflow.serving.PredictRequest")
PredictResponse = _cls("tensorflow.serving.PredictResponse")


def make_tensor_proto(values: np.ndarray, shape=None):
    """What `tf.make_tensor_proto(data, shape=data.shape)` produces for a numpy float32 array (model_server.py:35-36):
    dtype, tensor_shape and the raw little-endian bytes in `tensor_content`."""
    a = np.ascontiguousarray(values)
    if a.dtype != np.float32:
        raise TypeError(f"only float32 tensors cross this boundary, got {a.dtype}")
    p = TensorProto()
    p.dtype = DT_FLOAT
    for d in (shape if shape is not None else a.shape):
        p.tensor_shape.dim.add().size = int(d)
    p.tensor_content = a.tobytes()
    return p


def tensor_proto_to_ndarray(p) -> np.ndarray:
    shape = [int(d.size) for d in p.tensor_shape.dim]
    if p.dtype != DT_FLOAT:
        raise ValueError(f"expected DT_FLOAT (1), got dtype {p.dtype}")
    if p.tensor_content:
        a = np.frombuffer(p.tensor_content, dtype="<f4")
    else:
        a = np.asarray(p.float_val, dtype=np.float32)
        n = int(np.prod(shape)) if shape else a.size
        if a.size == 1 and n > 1:      # TF broadcast convention for a single repeated value
            a = np.full(n, a[0], np.float32)
    if int(np.prod(shape)) != a.size:
        raise ValueError(f"tensor_shape {shape} does not match {a.size} values")
    return a.reshape(shape)


class PredictionServicer:
    """Implements Predict on top of any `serving_default(input_8=ndarray) -> {'dense_7': ndarray}` callable
    (normally `engine.Model.serving_default`).  One call at a time reaches the model (a KdlModel is not re-entrant): either one
    request under a lock (the reference's configuration: TF-Serving without --enable_batching), or, with `batching`, one
    coalesced batch of concurrent requests from the batcher's scheduler thread (batching.py)."""

    def __init__(self, serving_default: Callable[..., Dict[str, np.ndarray]], input_size: int = 299, version: int = 1,
                 batching=None):
        self._fn = serving_default
        self._size = input_size
        self._version = version          # tf-serving.dockerfile:5 copies the model into version directory "1"
        self._lock = threading.Lock()
        self.batcher = None
        if batching is not None:
            from .batching import DynamicBatcher
            self.batcher = DynamicBatcher(self._run_batch, batching)

    def _run_batch(self, xs):
        """xs: list of (n_i, S, S, 3) float32 arrays -> list of (n_i, 10) arrays, one forward call for all of them."""
        x = xs[0] if len(xs) == 1 else np.concatenate(xs, axis=0)
        with self._lock:
            y = np.asarray(self._fn(**{INPUT_KEY: x})[OUTPUT_KEY], dtype=np.float32)
        out, o = [], 0
        for a in xs:
            out.append(y[o:o + a.shape[0]])
            o += a.shape[0]
        return out

    def close(self):
        if self.batcher is not None:
            self.batcher.close()

    def Predict(self, request, context):
        import grpc
        spec = request.model_spec
        if spec.name != MODEL_NAME:
            context.abort(grpc.StatusCode.NOT_FOUND, f"Servable not found for request: Latest({spec.name})")
        if spec.HasField("version") and spec.version.value != self._version:
            context.abort(grpc.StatusCode.NOT_FOUND, f"Servable not found for request: Specific({spec.name}, {spec.version.value})")
        sig = spec.signature_name or SIGNATURE_NAME
        if sig != SIGNATURE_NAME:
            context.abort(grpc.StatusCode.INVALID_ARGUMENT, f'Serving signature key "{sig}" not found.')
        if set(request.inputs.keys()) != {INPUT_KEY}:
            context.abort(grpc.StatusCode.INVALID_ARGUMENT,
                          f"input tensor alias not found in signature: expected exactly {{{INPUT_KEY}}}, got {sorted(request.inputs.keys())}")
        try:
            x = tensor_proto_to_ndarray(request.inputs[INPUT_KEY])
        except ValueError as e:
            context.abort(grpc.StatusCode.INVALID_ARGUMENT, str(e))
        S = self._size
        if x.ndim != 4 or x.shape[1:] != (S, S, 3):
            context.abort(grpc.StatusCode.INVALID_ARGUMENT,
                          f"input_8 must have shape (-1,{S},{S},3) (guide.md:224), got {tuple(x.shape)}")
        if self.batcher is not None and x.shape[0] > 0:
            from .batching import QueueFull
            try:
                y = self.batcher.submit(x, size=int(x.shape[0]))
            except QueueFull as e:
                context.abort(grpc.StatusCode.UNAVAILABLE, str(e))
        else:
            y = self._run_batch([x])[0]
        resp = PredictResponse()
        resp.model_spec.name = MODEL_NAME
        resp.model_spec.version.value = self._version
        resp.model_spec.signature_name = SIGNATURE_NAME
        out = resp.outputs[OUTPUT_KEY]
        out.dtype = DT_FLOAT
        for d in y.shape:
            out.tensor_shape.dim.add().size = int(d)
        out.float_val.extend(y.reshape(-1).tolist())   # the gateway reads float_val, not tensor_content (model_server.py:47)
        return resp


def make_server(serving_default: Callable[..., Dict[str, np.ndarray]], address: str = "127.0.0.1:8500", input_size: int = 299,
                max_workers: int = 4, max_message_mb: int = 512, batching=None):
    """Returns (grpc.Server, bound_port).  Port 8500 is the one the reference's Service exposes
    (tf-serving-clothing-model-service.yaml:9-10); pass ':0' for an ephemeral port.  `batching`: a
    batching.BatchingParameters turns on server-side request coalescing (give the server at least max_batch_size workers:
    every waiting request holds one)."""
    import grpc
    servicer = PredictionServicer(serving_default, input_size, batching=batching)
    if batching is not None:
        max_workers = max(max_workers, 2 * batching.max_batch_size)
    handler = grpc.method_handlers_generic_handler(SERVICE, {
        "Predict": grpc.unary_unary_rpc_method_handler(servicer.Predict, request_deserializer=PredictRequest.FromString,
                                                       response_serializer=PredictResponse.SerializeToString)})
    opts = [("grpc.max_receive_message_length", max_message_mb << 20), ("grpc.max_send_message_length", max_message_mb << 20)]
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=max_workers), options=opts)
    server.add_generic_rpc_handlers((handler,))
    port = server.add_insecure_port(address)      # insecure, like the reference's channel (model_server.py:15)
    server.kdl_servicer = servicer                # for servicer.batcher statistics and close()
    return server, port


class PredictionServiceStub:
    """Client side, shaped like `prediction_service_pb2_grpc.PredictionServiceStub` (model_server.py:16): `stub.Predict(req, timeout=20.0)`."""

    def __init__(self, channel):
        self.Predict = channel.unary_unary(f"/{SERVICE}/Predict", request_serializer=PredictRequest.SerializeToString,
                                           response_deserializer=PredictResponse.FromString)


def serve_forever(weights=None, address: str = "0.0.0.0:8500", device: int = 0, max_batch: int = 32,
                  model: Optional[object] = None, enable_batching: Optional[bool] = None):   # pragma: no cover - process entry point
    """`python -m kdl_b200.serving`: the container entry point that replaces `tensorflow/serving:2.3.0` (tf-serving.dockerfile).
    KDL_ENABLE_BATCHING=1 (or enable_batching=True) is the counterpart of TF-Serving's --enable_batching; KDL_BATCH_TIMEOUT_MICROS
    and KDL_MAX_BATCH_SIZE set the two parameters that matter."""
    import os
    from . import engine, weights as wmod
    from .batching import BatchingParameters
    if enable_batching is None:
        enable_batching = os.environ.get("KDL_ENABLE_BATCHING", "0") == "1"
    max_batch = int(os.environ.get("KDL_MAX_BATCH_SIZE", max_batch))
    if model is None:
        model = engine.Model(weights or wmod.load_from_env(), device=device, max_batch=max_batch)
    params = BatchingParameters(max_batch_size=max_batch,
                                batch_timeout_micros=int(os.environ.get("KDL_BATCH_TIMEOUT_MICROS", 1000))) if enable_batching else None
    server, port = make_server(model.serving_default, address, model.input_size, batching=params)
    server.start()
    print(f"PredictionService for '{MODEL_NAME}' listening on port {port}")
    server.wait_for_termination()


if __name__ == "__main__":   # pragma: no cover
    serve_forever()
