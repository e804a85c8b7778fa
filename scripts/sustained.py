"""Sustained throughput: T seconds of back-to-back batch-256 forward passes per tuning variant (device-resident input), with the
power / clock trace of nvidia-smi beside it.  After ~0.5 s the B200 sits at its 1000 W power limit and the step time is set by the
energy a step costs, not by its cycle count; bench.py's default 20 timed steps end before the limiter engages.
  python scripts/sustained.py [seconds] [json tuning] [json tuning] ..."""
import json, os, subprocess, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import kdl_b200  # noqa: F401
from kdl_b200 import engine, weights

T = float(sys.argv[1]) if len(sys.argv) > 1 else 6.0
variants = [json.loads(a) for a in sys.argv[2:]] or [{}]
W = weights.make_weights(1, "calibrated")
rng = np.random.default_rng(0)
pk = engine.PackedImages([rng.integers(0, 256, (299, 299, 3), dtype=np.uint8) for _ in range(256)])
d = pk.host.cuda()
out = torch.empty((256, 10), device="cuda")
for tun in variants:
    m = engine.Model(W, max_batch=256, precision="bf16", tuning=tun)
    for _ in range(5):
        m.predict_packed_device(d, pk, out)
    torch.cuda.synchronize()
    time.sleep(3.0)   # let the power average fall back
    smi = subprocess.Popen(["nvidia-smi", "--query-gpu=power.draw.instant,clocks.sm,clocks_event_reasons.active", "--format=csv,noheader,nounits",
                            "-lms", "100"], stdout=subprocess.PIPE, text=True)
    t0 = time.time()
    per_sec = []
    n = 0
    mark, nmark = t0, 0
    while time.time() - t0 < T:
        for _ in range(20):
            m.predict_packed_device(d, pk, out)
        torch.cuda.synchronize()
        n += 20
        now = time.time()
        if now - mark >= 1.0:
            per_sec.append((now - mark) / (n - nmark) * 1e3)
            mark, nmark = now, n
    total = (time.time() - t0) / n * 1e3
    smi.terminate()
    rows = [r.split(",") for r in smi.stdout.read().strip().splitlines()]
    busy = [(float(r[0]), float(r[1])) for r in rows if len(r) >= 2 and float(r[0]) > 300]
    pw = sorted(p for p, _ in busy)
    ck = sorted(c for _, c in busy)
    print(f"{json.dumps(tun):40s} {total:6.2f} ms/step over {T:.0f} s ({256e3 / total:7.0f} img/s); per second: "
          + " ".join(f"{x:.2f}" for x in per_sec) + f"; power median {pw[len(pw) // 2] if pw else 0:.0f} W, SM clock median {ck[len(ck) // 2] if ck else 0:.0f} MHz", flush=True)
    m.close()
