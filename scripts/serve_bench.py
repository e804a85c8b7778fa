"""Closed-loop load test of the PredictionService stand-in (serving.py) on one GPU: C client processes each send single-image
`Predict` requests (the reference gateway's request: 1x299x299x3 float32 in tensor_content) for T seconds, once without and once
with server-side batching.  Reports requests/s and client-side p50/p99.  Python gRPC + 1 MB protobuf messages cap this well below
what the engine does (bench.py measures that); the point is the ratio batching buys at the reference's own boundary."""
import argparse
import json
import time

import grpc
import numpy as np

import kdl_b200  # noqa: F401
from kdl_b200 import engine, serving, weights
from kdl_b200.batching import BatchingParameters


def _client(port, seconds, seed, q):
    """one client PROCESS (its own GIL): closed loop of single-image Predict calls"""
    rng = np.random.default_rng(seed)
    ch = grpc.insecure_channel(f"127.0.0.1:{port}", options=[("grpc.max_send_message_length", 64 << 20)])
    stub = serving.PredictionServiceStub(ch)
    X = (rng.integers(0, 256, (1, 299, 299, 3)).astype(np.float32) / np.float32(127.5) - np.float32(1.0))
    req = serving.PredictRequest()
    req.model_spec.name, req.model_spec.signature_name = serving.MODEL_NAME, serving.SIGNATURE_NAME
    req.inputs[serving.INPUT_KEY].CopyFrom(serving.make_tensor_proto(X))
    for _ in range(3):
        stub.Predict(req, timeout=20.0)
    lat = []
    stop = time.monotonic() + seconds
    while time.monotonic() < stop:
        t0 = time.perf_counter()
        stub.Predict(req, timeout=20.0)
        lat.append(time.perf_counter() - t0)
    ch.close()
    q.put(lat)


def run(model, batching, clients, seconds):
    import multiprocessing as mp
    server, port = serving.make_server(model.serving_default, "127.0.0.1:0", batching=batching, max_workers=2 * clients)
    server.start()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_client, args=(port, seconds, i, q)) for i in range(clients)]
    for p in ps:
        p.start()
    lat = [q.get(timeout=seconds + 120) for _ in ps]
    for p in ps:
        p.join(30)
    server.stop(0)
    sizes = list(server.kdl_servicer.batcher.batch_sizes) if batching else [1]
    server.kdl_servicer.close()
    a = np.sort(np.concatenate([np.asarray(l) for l in lat]))
    return {"batching": bool(batching), "clients": clients, "requests_per_s": round(len(a) / seconds, 1),
            "p50_ms": round(1e3 * a[len(a) // 2], 2), "p99_ms": round(1e3 * a[int(len(a) * 0.99)], 2),
            "mean_batch": round(float(np.mean(sizes)), 2)}


def run_inproc(model, batching, clients, seconds):
    """the gateway's in-process path: C request threads call Gateway.apply_model_array (decoded uint8 image -> {label: logit})"""
    import threading
    from kdl_b200 import gateway
    g = gateway.Gateway(model, batching=batching)
    rng = np.random.default_rng(1)
    imgs = [rng.integers(0, 256, (534, 400, 3), dtype=np.uint8) for _ in range(clients)]
    lat = [[] for _ in range(clients)]
    g.apply_model_array(imgs[0])
    stop = time.monotonic() + seconds

    def work(i):
        while time.monotonic() < stop:
            t0 = time.perf_counter()
            g.apply_model_array(imgs[i])
            lat[i].append(time.perf_counter() - t0)
    ts = [threading.Thread(target=work, args=(i,)) for i in range(clients)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    sizes = list(g._u8_batcher.batch_sizes) if batching else [1]
    g.close()
    a = np.sort(np.concatenate([np.asarray(l) for l in lat]))
    return {"path": "Gateway.apply_model_array (in-process, 534x400 uint8)", "batching": bool(batching), "clients": clients,
            "requests_per_s": round(len(a) / seconds, 1), "p50_ms": round(1e3 * a[len(a) // 2], 2),
            "p99_ms": round(1e3 * a[int(len(a) * 0.99)], 2), "mean_batch": round(float(np.mean(sizes)), 2)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--clients", type=int, nargs="+", default=[1, 8, 32])
    ap.add_argument("--seconds", type=float, default=4.0)
    ap.add_argument("--timeout-micros", type=int, default=1000)
    ap.add_argument("--inproc", action="store_true", help="measure Gateway.apply_model_array instead of the gRPC server")
    a = ap.parse_args()
    m = engine.Model(weights.make_weights(1, "calibrated"), max_batch=32)
    for c in a.clients:
        for b in (None, BatchingParameters(max_batch_size=32, batch_timeout_micros=a.timeout_micros)):
            print(json.dumps((run_inproc if a.inproc else run)(m, b, c, a.seconds)), flush=True)
    m.close()


if __name__ == "__main__":
    main()
