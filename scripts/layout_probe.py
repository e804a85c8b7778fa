"""Probe (instrumented build only): depthwise 3x3 at 19x19 on the pixel-major pitch-768 layout vs a chunk-major layout.

  make -C kubernetes-deep-learning_b200/csrc prof
  KDL_LIB=kubernetes-deep-learning_b200/csrc/libkdl_prof.so python scripts/layout_probe.py [batch]
"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: F401  (CUDA context + libcudart)
import kdl_b200  # noqa: F401
from kdl_b200 import _lib

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
lib = _lib.load()
out = (ctypes.c_float * 6)()
lib.kdl_layout_probe.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
for rnd in (0, 1, 0, 1):
    lib.kdl_layout_probe_data(rnd)
    print('operands:', 'pseudo-random' if rnd else 'constant input, zero weights')
    rc = lib.kdl_layout_probe(B, 40, out)
    assert rc == 0, rc
    useful = B * 361 * 728 * 2 * 2
    print(f"batch {B}: pixel-major pitch 768: {out[0]:.1f} us ({useful / out[0] / 1e3:.0f} GB/s)   "
          f"chunk-major 12 x 64: {out[1]:.1f} us ({useful / out[1] / 1e3:.0f} GB/s useful, {useful * 768 / 728 / out[1] / 1e3:.0f} GB/s moved)")
    fl = 2.0 * B * 361 * 728 * 728
    print(f"   GEMM 728x728 alone, alternating buffers: {out[2]:.1f} us ({fl / out[2] / 1e6:.0f} TFLOP/s); same buffers: {out[5]:.1f} us ({fl / out[5] / 1e6:.0f})")
    print(f"   dw -> GEMM chain: serpentine {out[3]:.1f} us, forwards {out[4]:.1f} us; sum of the stand-alone times {out[0] + out[2]:.1f} us")
