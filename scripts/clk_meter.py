"""Probe (instrumented build only): the SM clock the pointwise GEMM and the depthwise kernel actually run at, inside the batch-256
forward pass and in the stand-alone loops of kdl_layout_probe (clock64 against %globaltimer in CTA 0 of every launch).
  KDL_LIB=kubernetes-deep-learning_b200/csrc/libkdl_prof.so python scripts/clk_meter.py"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import kdl_b200  # noqa: F401
from kdl_b200 import _lib, engine, weights

lib = _lib.load()
buf = (ctypes.c_uint64 * 16)()


def read(tag):
    assert lib.kdl_clk_meter_read(buf, 1) == 0
    a = np.frombuffer(buf, dtype=np.uint64).reshape(4, 4).astype(np.float64)
    for slot, name in ((0, "pointwise GEMM"), (1, "depthwise"), (2, "conv2")):
        cyc, ns, n = a[slot][:3]
        if n:
            print(f"{tag:44s} {name:15s} {int(n):5d} launches  {ns / n / 1e3:8.1f} us each  {cyc / ns * 1e3:7.0f} MHz")


B = 256
W = weights.make_weights(1, "calibrated")
m = engine.Model(W, max_batch=B, precision="bf16")
rng = np.random.default_rng(0)
pk = engine.PackedImages([rng.integers(0, 256, (299, 299, 3), dtype=np.uint8) for _ in range(B)])
d = pk.host.cuda()
for _ in range(3):
    m.predict_packed_device(d, pk)
read("warm-up")
for _ in range(20):
    m.predict_packed_device(d, pk)
read("forward pass, batch 256, 20 steps")
m.close()
out = (ctypes.c_float * 6)()
lib.kdl_layout_probe.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
for rnd in (0, 1):
    lib.kdl_layout_probe_data(rnd)
    assert lib.kdl_layout_probe(B, 40, out) == 0
    read("stand-alone loops, " + ("pseudo-random" if rnd else "constant/zero") + " operands")
