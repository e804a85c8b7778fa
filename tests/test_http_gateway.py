"""HTTP shell of the gateway on CPU with a stub model: the /predict JSON contract of model_server.py:59-66."""
import json
import threading
import urllib.error
import urllib.request

import pytest

import kdl_b200  # noqa: F401
from kdl_b200 import gateway, http_gateway


def fake_apply_model(url):
    if "boom" in url:
        raise RuntimeError("cannot fetch")
    return {c: float(i) + len(url) for i, c in enumerate(gateway.labels)}


@pytest.fixture(scope="module")
def base():
    srv = http_gateway.make_http_server(fake_apply_model, "127.0.0.1", 0)
    t = threading.Thread(target=srv.serve_forever, daemon=True)
    t.start()
    yield f"http://127.0.0.1:{srv.server_address[1]}"
    srv.shutdown()


def post(url, obj):
    req = urllib.request.Request(url, data=json.dumps(obj).encode(), headers={"Content-Type": "application/json"})
    with urllib.request.urlopen(req, timeout=10) as r:
        return r.status, json.loads(r.read())


def test_predict_contract(base):
    # the body test.py:3-16 posts
    status, body = post(base + "/predict", {"url": "http://bit.ly/mlbookcamp-pants"})
    assert status == 200
    assert list(body) == gateway.labels                       # ten labels, reference order (model_server.py:21-32)
    assert body["dress"] == float(len("http://bit.ly/mlbookcamp-pants"))


def test_error_statuses(base):
    with pytest.raises(urllib.error.HTTPError) as e:
        post(base + "/predict", {"uri": "x"})
    assert e.value.code == 400
    with pytest.raises(urllib.error.HTTPError) as e:
        post(base + "/predict", {"url": "http://boom"})
    assert e.value.code == 500                                # uncaught exception -> 500, as under Flask
    with pytest.raises(urllib.error.HTTPError) as e:
        post(base + "/other", {"url": "x"})
    assert e.value.code == 404
    with pytest.raises(urllib.error.HTTPError) as e:
        urllib.request.urlopen(base + "/predict", timeout=10)
    assert e.value.code == 405


def test_health_route_for_probes(base):
    with urllib.request.urlopen(base + "/healthz", timeout=10) as r:
        assert r.status == 200 and json.loads(r.read()) == {"status": "ok"}


def test_weight_file_dispatch(tmp_path, monkeypatch):
    """KDL_WEIGHTS -> weights.load: Keras .h5 and .npz both give the bundle back; anything else is refused"""
    import numpy as np
    from kdl_b200 import keras_io, weights
    W = weights.make_weights(2, "keras_default", head_inner=0)
    p = tmp_path / "m.h5"
    keras_io.export_keras_h5(W, str(p))
    monkeypatch.setenv("KDL_WEIGHTS", str(p))
    got = weights.load_from_env()
    assert list(got) == list(W) and all(np.array_equal(got[k], W[k]) for k in W)
    with pytest.raises(ValueError, match="unknown weight file type"):
        weights.load(str(tmp_path / "m.ckpt"))


def test_entry_points_refuse_to_serve_without_weights(monkeypatch):
    """a missing KDL_WEIGHTS must fail loudly (TF-Serving does not start without its model either),
    unless random weights are asked for explicitly."""
    from kdl_b200 import weights
    monkeypatch.delenv("KDL_WEIGHTS", raising=False)
    monkeypatch.delenv("KDL_ALLOW_RANDOM_WEIGHTS", raising=False)
    with pytest.raises(RuntimeError, match="KDL_WEIGHTS is not set"):
        weights.load_from_env()
    monkeypatch.setenv("KDL_ALLOW_RANDOM_WEIGHTS", "1")
    assert "dense_7/kernel" in weights.load_from_env()
