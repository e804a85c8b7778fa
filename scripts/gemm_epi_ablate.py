"""Times the 92416 x 768 x 768 pointwise GEMM (single-CTA kernel) with and without a residual; run with KDL_LIB pointing at an
ablation build (-DKDL_ABLATE_EPI=1|2) to see what the epilogue costs."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, kdl_b200
from kdl_b200 import engine
M, N, K = 92416, 768, 768
A = torch.randn((M, K), device="cuda").to(torch.bfloat16); Wt = torch.randn((N, K), device="cuda").to(torch.bfloat16)
R = torch.randn((M, N), device="cuda").to(torch.bfloat16)
bias = torch.zeros(N, device="cuda"); out = torch.empty((M, N), dtype=torch.bfloat16, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for res in (None, R):
    ts = []
    for i in range(8):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); engine.gemm_bf16(A, Wt, bias, res, True, out=out); e1.record(); torch.cuda.synchronize()
        if i >= 2: ts.append(e0.elapsed_time(e1))
    t = sorted(ts)[len(ts) // 2]
    print(os.environ.get("KDL_LIB", "default"), "residual" if res is not None else "no residual", f"{t*1e3:.1f} us  {2*M*N*K/t*1e-9:.0f} TF")
