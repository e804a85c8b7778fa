"""Quick check of the CTA-pair fused middle-flow kernel (KdlTuning.fused_midflow = 2): bit-identical logits to the default path and
its step time.  python scripts/pair_check.py [batch]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import kdl_b200  # noqa: F401
from kdl_b200 import engine, weights

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
W = weights.make_weights(1, "calibrated")
rng = np.random.default_rng(0)
pk = engine.PackedImages([rng.integers(0, 256, (299, 299, 3), dtype=np.uint8) for _ in range(B)])
d = pk.host.cuda()
outs = {}
for name, tun in (("default", {}), ("fused_midflow=1", {"fused_midflow": 1}), ("fused_midflow=2 (pair)", {"fused_midflow": 2}),
                  ("fused_midflow=3 (trio)", {"fused_midflow": 3})):
    m = engine.Model(W, max_batch=B, precision="bf16", tuning=tun)
    out = torch.empty((B, 10), device="cuda")
    for _ in range(3):
        m.predict_packed_device(d, pk, out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(10):
        m.predict_packed_device(d, pk, out)
    e1.record(); torch.cuda.synchronize()
    outs[name] = out.cpu().numpy().copy()
    print(f"{name:24s} {e0.elapsed_time(e1) / 10:8.3f} ms/step  identical to default: {np.array_equal(outs[name], outs['default'])}", flush=True)
    m.close()
