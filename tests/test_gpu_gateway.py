"""The drop-in contract of model_server.py on the GPU: labels, names, apply_model-shaped results."""
import json
import os

import numpy as np
import pytest
import torch

from gpu_util import rel

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gw():
    import kdl_b200  # noqa: F401
    from kdl_b200 import engine, gateway, weights
    W = weights.make_weights(1, "calibrated")
    m = engine.Model(W, max_batch=8, precision="bf16")
    yield gateway.Gateway(m), W
    m.close()


def test_apply_model_contract_on_fixture(gw, golden_dir):
    from kdl_b200 import gateway
    from oracle import preprocess_oracle as po
    from oracle import xception_oracle as xo
    g, W = gw
    src = np.load(os.path.join(golden_dir, "image2_fixture.npz"))["src"]
    # the reference's call sequence (model_server.py:52-56) with the download step replaced by the fixture
    X = g.preprocessor.from_array(src)
    assert X.tobytes() == po.preprocess_pillow(src).tobytes()
    req = gateway.make_request(X)
    assert req["model_spec"] == {"name": "clothing-model", "signature_name": "serving_default"}
    assert list(req["inputs"]) == ["input_8"]
    res = g.Predict(req, timeout=20.0)
    assert list(res["outputs"]) == ["dense_7"] and res["outputs"]["dense_7"].shape == (1, 10)
    out = gateway.process_response(res)
    assert list(out) == ['dress', 'hat', 'longsleeve', 'outwear', 'pants', 'shirt', 'shoes', 'shorts', 'skirt', 't-shirt']
    assert all(isinstance(v, float) for v in out.values())
    json.dumps(out)   # jsonify-able, model_server.py:66
    want = xo.forward(W, po.preprocess_pillow(src), torch.float64)[0]
    got = np.array(list(out.values()))
    assert rel(got, want) <= 2e-2 and got.argmax() == want.argmax()
    # the in-process uint8 path (apply_model_array): same contract; it differs from the float-tensor path only by conv1's bf16
    # weight rounding (raw pixels go to the tensor cores there, KdlTuning.conv1_tc)
    fused = g.apply_model_array(src)
    assert list(fused) == list(out)
    fv = np.array(list(fused.values()), np.float32)
    assert rel(fv, want) <= 2e-2 and fv.argmax() == want.argmax()
    assert rel(fv, got) <= 1e-2


def test_unknown_model_or_signature_is_an_error(gw):
    from kdl_b200 import gateway
    g, _ = gw
    X = np.zeros((1, 299, 299, 3), np.float32)
    req = gateway.make_request(X)
    req["model_spec"]["name"] = "other"
    with pytest.raises(LookupError):
        g.Predict(req)
    req = gateway.make_request(X)
    req["model_spec"]["signature_name"] = "predict"
    with pytest.raises(LookupError):
        g.Predict(req)


def test_process_response_keeps_reference_batch_semantics(gw):
    """model_server.py:47-48 zips 10 labels against the flat float list: a batch returns image 0 only."""
    from kdl_b200 import gateway
    r = gateway.process_response({"outputs": {"dense_7": np.arange(30, dtype=np.float32).reshape(3, 10)}})
    assert list(r.values()) == [float(i) for i in range(10)]


def test_predict_batch_mixed_sizes(gw):
    from oracle import preprocess_oracle as po
    from oracle import xception_oracle as xo
    g, W = gw
    rng = np.random.default_rng(2)
    imgs = [rng.integers(0, 256, (int(rng.integers(64, 2049)), int(rng.integers(64, 2049)), 3), dtype=np.uint8)
            for _ in range(6)]
    res = g.predict_batch(imgs)
    x = np.concatenate([po.preprocess_pillow(im) for im in imgs])
    want = xo.forward(W, x, torch.float64)
    for i, r in enumerate(res):
        got = np.array(list(r.values()))
        assert rel(got, want[i]) <= 2e-2 and got.argmax() == want[i].argmax()


def test_replica_set_matches_single_model():
    """In-process replicas over every visible GPU give the same logits, in input order, as one model."""
    import kdl_b200  # noqa: F401
    from kdl_b200 import engine, replicas, weights
    W = weights.make_weights(1, "calibrated")
    rng = np.random.default_rng(31)
    imgs = [rng.integers(0, 256, (int(rng.integers(64, 600)), int(rng.integers(64, 600)), 3), dtype=np.uint8) for _ in range(7)]
    single = engine.Model(W, max_batch=8, precision="bf16")
    want = single.predict_uint8(imgs)
    single.close()
    rs = replicas.ReplicaSet(W, max_batch=8, precision="bf16")
    try:
        got = rs.predict_uint8(imgs)
        assert len(rs.models) == torch.cuda.device_count()
    finally:
        rs.close()
    np.testing.assert_array_equal(got, want)


def test_unmodified_gateway_flow_over_grpc_and_http(gw, tmp_path):
    """End to end through both shells of SURVEY.md §8f: HTTP POST /predict -> apply_model (file:// URL) -> engine, and the
    model_server.py call sequence against the PredictionService stand-in over a real gRPC channel."""
    import json
    import threading
    import urllib.request
    grpc = pytest.importorskip("grpc")
    from PIL import Image
    from kdl_b200 import gateway, http_gateway, serving
    from oracle import preprocess_oracle as po
    from oracle import xception_oracle as xo
    g, W = gw
    rng = np.random.default_rng(8)
    img = rng.integers(0, 256, (333, 444, 3), dtype=np.uint8)
    path = tmp_path / "pic.png"
    Image.fromarray(img).save(path)
    want = xo.forward(W, po.preprocess_pillow(img), torch.float64)[0]

    # --- gRPC: what model_server.py does with TF_SERVING_HOST pointing at this server
    server, port = serving.make_server(g.model.serving_default, "127.0.0.1:0")
    server.start()
    try:
        channel = grpc.insecure_channel(f"127.0.0.1:{port}", options=[("grpc.max_send_message_length", 64 << 20)])
        stub = serving.PredictionServiceStub(channel)
        X = g.preprocessor.from_path(str(path))
        pb_request = serving.PredictRequest()
        pb_request.model_spec.name = 'clothing-model'
        pb_request.model_spec.signature_name = 'serving_default'
        pb_request.inputs['input_8'].CopyFrom(serving.make_tensor_proto(X, shape=X.shape))
        pb_result = stub.Predict(pb_request, timeout=20.0)
        pred = np.array(pb_result.outputs['dense_7'].float_val)
        assert rel(pred, want) <= 2e-2 and pred.argmax() == want.argmax()
        channel.close()
    finally:
        server.stop(0)

    # --- HTTP: POST /predict {"url": ...}
    srv = http_gateway.make_http_server(g.apply_model, "127.0.0.1", 0)
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    try:
        req = urllib.request.Request(f"http://127.0.0.1:{srv.server_address[1]}/predict",
                                     data=json.dumps({"url": path.as_uri()}).encode(), headers={"Content-Type": "application/json"})
        with urllib.request.urlopen(req, timeout=30) as r:
            body = json.loads(r.read())
        assert list(body) == gateway.labels
        got = np.array(list(body.values()))
        assert rel(got, want) <= 2e-2 and got.argmax() == want.argmax()
        assert rel(got, pred) <= 1e-2      # same engine; the uint8 route differs from the float-tensor (gRPC) route by conv1's weight rounding
    finally:
        srv.shutdown()


def test_batched_gateway_gives_each_concurrent_request_its_own_unbatched_answer(gw):
    """Gateway(batching=...) (request coalescing): 20 threads call apply_model_array at once with images of
    different sizes; each must get bit-identical logits to the same image sent alone, and the model must have seen batches."""
    import threading
    from kdl_b200 import gateway
    from kdl_b200.batching import BatchingParameters
    g, _ = gw
    rng = np.random.default_rng(21)
    imgs = [rng.integers(0, 256, (int(rng.integers(64, 700)), int(rng.integers(64, 700)), 3), dtype=np.uint8) for _ in range(20)]
    alone = [g.apply_model_array(im) for im in imgs]
    gb = gateway.Gateway(g.model, batching=BatchingParameters(max_batch_size=min(8, g.model.max_batch), batch_timeout_micros=100000))
    try:
        got = [None] * len(imgs)

        def work(i):
            got[i] = gb.apply_model_array(imgs[i])
        ts = [threading.Thread(target=work, args=(i,)) for i in range(len(imgs))]
        for t in ts:
            t.start()
        for t in ts:
            t.join(60)
        assert got == alone
        assert max(gb._u8_batcher.batch_sizes) > 1 and gb._u8_batcher.images_run == len(imgs)
        # the float path (Predict) is coalesced too: identical to the same request sent through the unbatched gateway
        X = g.preprocessor.from_array(imgs[0])
        y = gb.Predict(gateway.make_request(X), timeout=20.0)
        assert gateway.process_response(y) == gateway.process_response(g.Predict(gateway.make_request(X), timeout=20.0))
    finally:
        gb.close()
