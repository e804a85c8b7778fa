"""92416 x 768 x 768 pointwise GEMM: single-CTA kernel vs CTA-pair (cta_group::2) vs 256x256-per-CTA, L2 flushed between runs."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, kdl_b200
from kdl_b200 import engine
M, N, K = 92416, 768, 768
A = torch.randn((M, K), device="cuda").to(torch.bfloat16); Wt = torch.randn((N, K), device="cuda").to(torch.bfloat16)
bias = torch.zeros(N, device="cuda"); out = torch.empty((M, N), dtype=torch.bfloat16, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for name, kw in (("single-CTA 128x256", dict(one_cta=True)), ("CTA pair 256x256", dict()), ("256x256 per CTA", dict(m256=True))):
    ts = []
    for i in range(8):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); engine.gemm_bf16(A, Wt, bias, None, True, out=out, **kw); e1.record(); torch.cuda.synchronize()
        if i >= 2: ts.append(e0.elapsed_time(e1))
    t = sorted(ts)[len(ts) // 2]
    print(f"{name:<22} {t*1e3:7.1f} us  {2*M*N*K/t*1e-9:6.0f} TFLOP/s")
