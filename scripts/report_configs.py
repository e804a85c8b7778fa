#!/usr/bin/env python
"""Numbers for the configurations of BASELINE.json / SURVEY.md §8d that bench.py's single line does not carry:
accuracy of the bf16 path over many images, batch-64 and 150x150 throughput, B=1024 as successive calls, mixed-size
end-to-end latency.  Writes one JSON document (default gpurun_out/configs.json)."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import kdl_b200  # noqa: F401
from kdl_b200 import engine, weights
from oracle import preprocess_oracle as po
from oracle import xception_oracle as xo


def throughput(model, B, S, steps=10, warm=3):
    rng = np.random.default_rng(5)
    packs = [engine.PackedImages([rng.integers(0, 256, (S, S, 3), dtype=np.uint8) for _ in range(B)]) for _ in range(2)]
    d = [p.host.cuda() for p in packs]
    out = torch.empty((B, 10), device="cuda")
    for i in range(warm):
        model.predict_packed_device(d[i % 2], packs[i % 2], out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for i in range(steps):
        model.predict_packed_device(d[i % 2], packs[i % 2], out)
    e1.record()
    torch.cuda.synchronize()
    return B * steps / (e0.elapsed_time(e1) * 1e-3)


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "configs.json")
    n_acc = int(os.environ.get("KDL_ACC_IMAGES", "96"))
    rep = {}
    W = {init: weights.make_weights(1, init) for init in ("calibrated", "keras_default")}

    # ---- config 1: accuracy of the bf16 path @299 over n_acc images, both inits
    rng = np.random.default_rng(0)
    u8 = [rng.integers(0, 256, (299, 299, 3), dtype=np.uint8) for _ in range(n_acc)]
    x = np.concatenate([po.preprocess_np(im) for im in u8])
    acc = {}
    for init in W:
        want = np.concatenate([xo.forward(W[init], x[i:i + 16], torch.float64) for i in range(0, n_acc, 16)])
        m = engine.Model(W[init], max_batch=64, precision="bf16")
        got = m.predict_uint8(u8)
        m.close()
        errs = np.array([xo.rel_err(got[i], want[i]) for i in range(n_acc)])
        acc[init] = {"images": n_acc, "rel_err_mean": float(errs.mean()), "rel_err_p50": float(np.median(errs)),
                     "rel_err_p95": float(np.percentile(errs, 95)), "rel_err_max": float(errs.max()),
                     "frac_within_2e-2": float((errs <= 2e-2).mean()),
                     "top1_agreement": float((got.argmax(1) == want.argmax(1)).mean())}
        mf = engine.Model(W[init], max_batch=16, precision="f32")
        gotf = mf.serving_default(input_8=x[:16])["dense_7"]
        mf.close()
        ef = [xo.rel_err(gotf[i], want[i]) for i in range(16)]
        acc[init]["fp32_path_rel_err_max_16img"] = float(max(ef))
    rep["accuracy_bf16_299"] = acc

    # ---- throughputs
    m64 = engine.Model(W["calibrated"], max_batch=64, precision="bf16")
    rep["bf16_b64_299_img_s"] = throughput(m64, 64, 299)
    m64.close()
    m150 = engine.Model(W["calibrated"], max_batch=256, input_size=150, precision="bf16")
    rep["bf16_b256_150_img_s"] = throughput(m150, 256, 150)
    m150.close()
    m256 = engine.Model(W["calibrated"], max_batch=256, precision="bf16")
    t = throughput(m256, 256, 299)
    rep["bf16_b256_299_img_s"] = t
    rep["b1024_on_1gpu_as_4x256_img_s"] = t

    # ---- config 4: mixed sizes end to end (host pixels -> host logits), batches 1..32
    rng = np.random.default_rng(2)
    lat = {}
    for B in (1, 2, 4, 8, 16, 32):
        calls = []
        for _ in range(6):
            imgs = [rng.integers(0, 256, (int(rng.integers(64, 2049)), int(rng.integers(64, 2049)), 3), dtype=np.uint8)
                    for _ in range(B)]
            calls.append(engine.PackedImages(imgs))
        h = np.empty((B, 10), np.float32)
        for c in calls:
            m256.predict_packed_host(c, h)
        ts = []
        for i in range(120):
            c = calls[i % len(calls)]
            t0 = time.perf_counter()
            m256.predict_packed_host(c, h)
            ts.append((time.perf_counter() - t0) * 1e3)
        ts.sort()
        lat[str(B)] = {"p50_ms": ts[len(ts) // 2], "p99_ms": ts[int(len(ts) * 0.99)],
                       "mean_src_mb": float(np.mean([c.total_bytes for c in calls]) / 1e6)}
    rep["mixed_size_e2e_latency"] = lat
    m256.close()
    json.dump(rep, open(out_path, "w"), indent=1)
    print(json.dumps(rep, indent=1))


if __name__ == "__main__":
    main()
