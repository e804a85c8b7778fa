"""Does the pointwise GEMM run faster when its A operand is already in L2?  M = 30720 rows (A = 47 MB, out = 47 MB: both fit the
126 MB L2) timed back to back (A resident) and with an L2 flush in between (A from HBM)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, kdl_b200
from kdl_b200 import engine
N = K = 768
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for M in (30720, 92416):
    A = torch.randn((M, K), device="cuda").to(torch.bfloat16); Wt = torch.randn((N, K), device="cuda").to(torch.bfloat16)
    bias = torch.zeros(N, device="cuda"); out = torch.empty((M, N), dtype=torch.bfloat16, device="cuda")
    for mode in ("flushed", "back-to-back"):
        ts = []
        for i in range(10):
            if mode == "flushed":
                flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); engine.gemm_bf16(A, Wt, bias, None, True, out=out); e1.record(); torch.cuda.synchronize()
            if i >= 3: ts.append(e0.elapsed_time(e1))
        t = sorted(ts)[len(ts) // 2]
        print(f"M={M:6d} {mode:<13} {t*1e3:7.1f} us  {2*M*N*K/t*1e-9:6.0f} TFLOP/s")
