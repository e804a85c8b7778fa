"""Probe: where the fused middle-flow kernel's roles spend their cycles.  Needs the instrumented build
(make -C kubernetes-deep-learning_b200/csrc prof) and KDL_LIB pointing at it:

  KDL_LIB=kubernetes-deep-learning_b200/csrc/libkdl_prof.so python scripts/mid_prof.py [batch]

Runs a batch-256 forward (so the last k_sepconv_mid launch is a 728 -> 728 layer at 19 x 19) and prints, averaged over the
CTAs, the cycles each role waited on each barrier and the cycles it worked."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import kdl_b200  # noqa: F401
from kdl_b200 import _lib, engine, weights

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
W = weights.make_weights(1, "calibrated")
m = engine.Model(W, max_batch=B, precision="bf16", use_graphs=False, tuning={"fused_midflow": int(os.environ.get("KDL_MIDFLOW", "1"))})
rng = np.random.default_rng(0)
pk = engine.PackedImages([rng.integers(0, 256, (299, 299, 3), dtype=np.uint8) for _ in range(B)])
d = pk.host.cuda()
for _ in range(3):
    m.predict_packed_device(d, pk)
torch.cuda.synchronize()
lib = _lib.load()
n = 148
buf = (ctypes.c_uint64 * (32 * n))()
lib.kdl_mid_prof_read.argtypes = [ctypes.c_void_p, ctypes.c_int]
assert lib.kdl_mid_prof_read(buf, n) == 0
a = np.frombuffer(buf, dtype=np.uint64).reshape(n, 32).astype(np.float64)
names = ["prod wait x_empty", "prod wait b_empty", "prod issue", "mma wait tempty", "mma wait a_full", "mma wait b_full", "mma issue",
         "epi wait tfull", "epi work", "dw6 wait x_full", "dw6 wait a_empty", "dw6 compute", "dw15 wait x_full", "dw15 wait a_empty",
         "dw15 compute"]
mode = os.environ.get("KDL_MIDFLOW", "1")
act = a[:, 6] > 0                      # CTAs that ran (a cluster launch may use fewer than 148)
a = a[act]
tiles = -(-(B * 361) // 128)
items = tiles * 2 / 148 if mode == "1" else tiles / (act.sum() / (2 if mode == "2" else 3))
print(f"batch {B}: {int(act.sum())} CTAs ran, {items:.2f} items per CTA (last launch: block13_sepconv1 728->728 @19x19)")
for i, nm in enumerate(names):
    print(f"{nm:22s} mean {a[:, i].mean():10.0f} clk   per item {a[:, i].mean() / items:8.0f}   min {a[:, i].min():9.0f} max {a[:, i].max():9.0f}")
m.close()
