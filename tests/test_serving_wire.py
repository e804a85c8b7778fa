"""The PredictionService stand-in speaks the wire format model_server.py expects.  CPU only: the model behind
the server is a stub; the GPU variant lives in test_gpu_gateway.py."""
import numpy as np
import pytest

import kdl_b200  # noqa: F401
from kdl_b200 import gateway, serving

grpc = pytest.importorskip("grpc")


def fake_serving_default(**inputs):
    x = inputs["input_8"]
    # deterministic logits that depend on the pixels: per-image mean + class index
    return {"dense_7": (x.reshape(x.shape[0], -1).mean(1, keepdims=True) + np.arange(10, dtype=np.float32)[None]).astype(np.float32)}


@pytest.fixture(scope="module")
def stub():
    server, port = serving.make_server(fake_serving_default, "127.0.0.1:0")
    server.start()
    channel = grpc.insecure_channel(f"127.0.0.1:{port}", options=[("grpc.max_send_message_length", 64 << 20),
                                                                  ("grpc.max_receive_message_length", 64 << 20)])
    yield serving.PredictionServiceStub(channel)
    channel.close()
    server.stop(0)


def make_request(X):
    """model_server.py:38-43, with serving.make_tensor_proto standing in for tf.make_tensor_proto."""
    pb_request = serving.PredictRequest()
    pb_request.model_spec.name = 'clothing-model'
    pb_request.model_spec.signature_name = 'serving_default'
    pb_request.inputs['input_8'].CopyFrom(serving.make_tensor_proto(X, shape=X.shape))
    return pb_request


def test_known_answer_wire_sizes():
    """SURVEY.md §9 probe P7: the request TensorProto for (1,299,299,3) float32 is 1 072 838 bytes, a 10-float response 54."""
    X = np.zeros((1, 299, 299, 3), np.float32)
    assert len(serving.make_tensor_proto(X).SerializeToString()) == 1_072_838
    t = serving.TensorProto()
    t.dtype = serving.DT_FLOAT
    t.tensor_shape.dim.add().size = 1
    t.tensor_shape.dim.add().size = 10
    t.float_val.extend([float(i) + 0.5 for i in range(10)])
    assert len(t.SerializeToString()) == 54


def test_tensorproto_is_interchangeable_with_the_real_tensorflow_definition():
    tb = pytest.importorskip("tensorboard.compat.proto.tensor_pb2")
    rng = np.random.default_rng(0)
    X = rng.standard_normal((2, 3, 4, 3)).astype(np.float32)
    ours = serving.make_tensor_proto(X)
    theirs = tb.TensorProto.FromString(ours.SerializeToString())          # parse ours with TF's own descriptor
    assert theirs.dtype == 1 and [d.size for d in theirs.tensor_shape.dim] == [2, 3, 4, 3]
    assert theirs.tensor_content == X.tobytes()
    back = serving.TensorProto.FromString(theirs.SerializeToString())
    np.testing.assert_array_equal(serving.tensor_proto_to_ndarray(back), X)
    theirs2 = tb.TensorProto(dtype=1, float_val=[1.0, 2.5, -3.0])
    theirs2.tensor_shape.dim.add().size = 3
    np.testing.assert_array_equal(serving.tensor_proto_to_ndarray(serving.TensorProto.FromString(theirs2.SerializeToString())),
                                  np.array([1.0, 2.5, -3.0], np.float32))


def test_predict_round_trip_like_the_gateway(stub):
    rng = np.random.default_rng(1)
    X = rng.uniform(-1, 1, (1, 299, 299, 3)).astype(np.float32)
    pb_result = stub.Predict(make_request(X), timeout=20.0)                 # model_server.py:55
    pred = pb_result.outputs['dense_7'].float_val                          # model_server.py:47
    result = {c: p for c, p in zip(gateway.labels, pred)}                  # model_server.py:48
    want = fake_serving_default(input_8=X)["dense_7"][0]
    assert list(result) == gateway.labels
    np.testing.assert_allclose(list(result.values()), want, rtol=0, atol=0)
    assert [d.size for d in pb_result.outputs['dense_7'].tensor_shape.dim] == [1, 10]
    assert pb_result.model_spec.name == 'clothing-model' and pb_result.model_spec.version.value == 1


def test_dynamic_batch_dimension(stub):
    X = np.zeros((3, 299, 299, 3), np.float32)
    X[1] += 0.5
    r = stub.Predict(make_request(X), timeout=20.0)
    assert len(r.outputs['dense_7'].float_val) == 30
    assert [d.size for d in r.outputs['dense_7'].tensor_shape.dim] == [3, 10]


def test_errors_match_tf_serving_status_codes(stub):
    X = np.zeros((1, 299, 299, 3), np.float32)
    req = make_request(X)
    req.model_spec.name = "other-model"
    with pytest.raises(grpc.RpcError) as e:
        stub.Predict(req, timeout=5.0)
    assert e.value.code() == grpc.StatusCode.NOT_FOUND
    req = make_request(X)
    req.model_spec.signature_name = "classify"
    with pytest.raises(grpc.RpcError) as e:
        stub.Predict(req, timeout=5.0)
    assert e.value.code() == grpc.StatusCode.INVALID_ARGUMENT
    req = serving.PredictRequest()
    req.model_spec.name = 'clothing-model'
    req.inputs['images'].CopyFrom(serving.make_tensor_proto(X))
    with pytest.raises(grpc.RpcError) as e:
        stub.Predict(req, timeout=5.0)
    assert e.value.code() == grpc.StatusCode.INVALID_ARGUMENT
    with pytest.raises(grpc.RpcError) as e:
        stub.Predict(make_request(np.zeros((1, 150, 150, 3), np.float32)), timeout=5.0)
    assert e.value.code() == grpc.StatusCode.INVALID_ARGUMENT


def test_server_side_batching_coalesces_concurrent_requests_and_routes_rows_back():
    """make_server(batching=...): 24 concurrent single-image Predict calls reach the model in fewer, larger calls, and every
    client still receives the logits of its own image."""
    import threading
    from kdl_b200.batching import BatchingParameters
    calls = []

    def counting(**inputs):
        calls.append(inputs["input_8"].shape[0])
        return fake_serving_default(**inputs)
    server, port = serving.make_server(counting, "127.0.0.1:0", input_size=8,
                                       batching=BatchingParameters(max_batch_size=8, batch_timeout_micros=50000))
    server.start()
    try:
        channel = grpc.insecure_channel(f"127.0.0.1:{port}")
        stub = serving.PredictionServiceStub(channel)
        rng = np.random.default_rng(5)
        Xs = [rng.standard_normal((1, 8, 8, 3)).astype(np.float32) for _ in range(24)]
        got = [None] * len(Xs)

        def work(i):
            got[i] = np.array(stub.Predict(make_request(Xs[i]), timeout=20.0).outputs['dense_7'].float_val, np.float32)
        ts = [threading.Thread(target=work, args=(i,)) for i in range(len(Xs))]
        for t in ts:
            t.start()
        for t in ts:
            t.join(30)
        for X, g in zip(Xs, got):
            np.testing.assert_array_equal(g, fake_serving_default(input_8=X)["dense_7"][0])
        assert sum(calls) == 24 and max(calls) <= 8 and len(calls) < 24
        # a multi-image request keeps its rows together
        X3 = np.concatenate(Xs[:3], 0)
        r = stub.Predict(make_request(X3), timeout=20.0).outputs['dense_7']
        assert [d.size for d in r.tensor_shape.dim] == [3, 10]
        np.testing.assert_array_equal(np.array(r.float_val, np.float32).reshape(3, 10), fake_serving_default(input_8=X3)["dense_7"])
        channel.close()
    finally:
        server.stop(0)
        server.kdl_servicer.close()
