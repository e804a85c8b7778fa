"""Probe (instrumented build only): where the stand-alone depthwise kernel's time goes at 19x19x728, batch 256.
  KDL_LIB=kubernetes-deep-learning_b200/csrc/libkdl_prof.so python scripts/dw_ablate.py"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: F401
import kdl_b200  # noqa: F401
from kdl_b200 import _lib
lib = _lib.load()
out = (ctypes.c_float * 6)()
lib.kdl_layout_probe.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
lib.kdl_layout_probe_data(1)
names = {0: "full kernel", 1: "TMA loads only (consumers wait and release)", 2: "loads + stores, no arithmetic", 3: "loads + arithmetic, no stores",
         4: "stores only (no loads)", 5: "loads + 16-byte stores (4 pixels per warp)", 6: "16-byte stores only"}
for rep in range(2):
    for mode in (0, 1, 2, 3, 4, 5, 6):
        lib.kdl_dw_ablate(mode)
        assert lib.kdl_layout_probe(256, 30, out) == 0
        print(f"{names[mode]:48s} pixel-major {out[0]:6.1f} us   chunk-major {out[1]:6.1f} us")
lib.kdl_dw_ablate(0)
