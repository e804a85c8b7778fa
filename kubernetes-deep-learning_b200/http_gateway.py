"""HTTP shell of the gateway: the `/predict` route of model_server.py:59-70 on the standard library
(flask / gunicorn are not installed here).  Same contract: `POST /predict` with JSON `{"url": ...}` answers a JSON object with the
ten clothing labels in the reference's order and raw logits as values; the dev server listens on 0.0.0.0:9696 (model_server.py:70,
serving-gateway-deployment.yaml).  What sits behind the route is any `apply_model(url) -> dict` callable -- normally
`gateway.Gateway(model).apply_model`, i.e. download + decode on the CPU and everything else on the GPU."""
import json
import threading
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Callable, Dict


def make_handler(apply_model: Callable[[str], Dict[str, float]], concurrent: bool = False):
    """concurrent=False: one request at a time reaches apply_model, like the single gunicorn sync worker (gateway.dockerfile:17).
    concurrent=True: request threads call apply_model in parallel -- for a Gateway built with `batching`, which does its own
    locking and coalesces the concurrent requests into one forward call."""
    import contextlib
    lock = contextlib.nullcontext() if concurrent else threading.Lock()

    class Handler(BaseHTTPRequestHandler):
        server_version = "kdl-gateway/0.1"

        def _send(self, code, obj):
            body = json.dumps(obj).encode()
            self.send_response(code)
            self.send_header("Content-Type", "application/json")
            self.send_header("Content-Length", str(len(body)))
            self.end_headers()
            self.wfile.write(body)

        def do_POST(self):
            if self.path.rstrip("/") != "/predict":
                return self._send(404, {"error": "not found"})
            try:
                n = int(self.headers.get("Content-Length", "0"))
                url = json.loads(self.rfile.read(n))       # request.get_json()  (model_server.py:64)
                target = url["url"]
            except (ValueError, KeyError, TypeError):
                return self._send(400, {"error": "expected a JSON body like {\"url\": \"http://...\"}"})
            try:
                with lock:
                    result = apply_model(target)           # model_server.py:65
            except Exception as e:                          # Flask answers 500 for an uncaught exception
                return self._send(500, {"error": f"{type(e).__name__}: {e}"})
            self._send(200, result)                        # jsonify(result)  (model_server.py:66)

        def do_GET(self):
            if self.path.rstrip("/") == "/healthz":    # for readiness probes (deploy/); the reference has no health route
                return self._send(200, {"status": "ok"})
            self._send(405, {"error": "method not allowed; POST /predict"})

        def log_message(self, fmt, *args):   # PYTHONUNBUFFERED-style stdout logging is the reference's only logging
            pass

    return Handler


def make_http_server(apply_model: Callable[[str], Dict[str, float]], host: str = "0.0.0.0", port: int = 9696,
                     concurrent: bool = False) -> ThreadingHTTPServer:
    return ThreadingHTTPServer((host, port), make_handler(apply_model, concurrent))


def main():   # pragma: no cover - process entry point
    import os
    from . import engine, gateway, weights
    from .batching import BatchingParameters
    model = engine.Model(weights.load_from_env(), max_batch=32)
    batching = os.environ.get("KDL_ENABLE_BATCHING", "0") == "1"
    params = BatchingParameters(32, int(os.environ.get("KDL_BATCH_TIMEOUT_MICROS", 1000))) if batching else None
    srv = make_http_server(gateway.Gateway(model, batching=params).apply_model, concurrent=batching)
    print("gateway listening on 0.0.0.0:9696")
    srv.serve_forever()


if __name__ == "__main__":   # pragma: no cover
    main()
