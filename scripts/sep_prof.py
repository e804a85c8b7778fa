"""Probe: where k_sepconv_tc's roles spend their cycles, per launch (blocks 2-3).  Needs the instrumented build:
  make -C kubernetes-deep-learning_b200/csrc prof
  KDL_LIB=kubernetes-deep-learning_b200/csrc/libkdl_prof.so python scripts/sep_prof.py [batch]"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import kdl_b200  # noqa: F401
from kdl_b200 import _lib, engine, weights

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
m = engine.Model(weights.make_weights(1, "calibrated"), max_batch=B, precision="bf16", use_graphs=False)
rng = np.random.default_rng(0)
pk = engine.PackedImages([rng.integers(0, 256, (299, 299, 3), dtype=np.uint8) for _ in range(B)])
d = pk.host.cuda()
for _ in range(3):
    m.predict_packed_device(d, pk)
torch.cuda.synchronize()
lib = _lib.load()
buf = (ctypes.c_uint64 * (4 * 160 * 16))()
lib.kdl_sep_prof_read.argtypes = [ctypes.c_void_p]
assert lib.kdl_sep_prof_read(buf) == 0
a = np.frombuffer(buf, dtype=np.uint64).reshape(4, 160, 16).astype(np.float64)[:, :148]
names = ["prod wait x_empty", "prod wait b_empty", "prod issue", "mma wait tempty", "mma wait a_full", "mma wait b_full", "mma issue",
         "epi wait tfull", "epi math+sts", "dw wait x_full", "dw wait a_empty", "dw compute", "epi wait store-read", "epi wait tmem ld",
         "epi fence+store issue"]
layers = [("block2_sepconv1 64->128 @147", 147, 1), ("block2_sepconv2 128->128 @147", 147, 2), ("block3_sepconv1 128->256 @74", 74, 2),
          ("block3_sepconv2 256->256 @74", 74, 4)]
for s, (nm, H, kch) in enumerate(layers):
    tiles = B * (-(-H // 16)) * (-(-H // 8)) / 148
    tot = a[s, :, 9:12].sum(1).mean()
    print(f"{nm}: {tiles:.1f} tiles per CTA x {kch} chunks; dw-warp timeline {tot:.0f} clk = {tot / tiles:.0f} per tile")
    for i, n in enumerate(names):
        print(f"   {n:20s} {a[s, :, i].mean():10.0f} clk   per tile {a[s, :, i].mean() / tiles:7.0f}   per chunk {a[s, :, i].mean() / tiles / kch:7.0f}")
m.close()
