"""HTTP shell of the gateway: the `/predict` route of model_server.py:59-70 on the standard library
(flask / gunicorn are not installed here).  Same contract: `POST /predict` with JSON `{"url": ...}` answers a JSON object with the
ten clothing labels in the reference's order and raw logits as values; the dev server listens on 0.0.0.0:9696 (model_server.py:70,
serving-gateway-deployment.yaml).  What sits behind the route is any `apply_model(url) -> dict` callable -- normally
`gateway.Gateway(model).apply_model`, i.e. download + decode on the CPU and everything else on the GPU."""
import json
import threading
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Callable, Dict


def make_handler(apply_model: Callable[[str], Dict[str, float]]):
    lock = threading.Lock()   # one request at a time reaches the model, like the single gunicorn sync worker (gateway.dockerfile:17)

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
            self._send(405, {"error": "method not allowed; POST /predict"})

        def log_message(self, fmt, *args):   # PYTHONUNBUFFERED-style stdout logging is the reference's only logging
            pass

    return Handler


def make_http_server(apply_model: Callable[[str], Dict[str, float]], host: str = "0.0.0.0", port: int = 9696) -> ThreadingHTTPServer:
    return ThreadingHTTPServer((host, port), make_handler(apply_model))


def main():   # pragma: no cover - process entry point
    from . import engine, gateway, weights
    model = engine.Model(weights.make_weights(1, "calibrated"), max_batch=32)
    srv = make_http_server(gateway.Gateway(model).apply_model)
    print("gateway listening on 0.0.0.0:9696")
    srv.serve_forever()


if __name__ == "__main__":   # pragma: no cover
    main()
