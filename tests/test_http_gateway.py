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
