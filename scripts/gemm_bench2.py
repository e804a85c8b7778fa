"""tcgen05 GEMM variants vs cuBLAS (torch.matmul, yardstick only) on square-ish and network shapes."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, kdl_b200
from kdl_b200 import engine

def timeit(fn, iters=10):
    for _ in range(3): fn()
    torch.cuda.synchronize(); e0 = torch.cuda.Event(True); e1 = torch.cuda.Event(True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters

for M, N, K in [(92416, 768, 768), (92416, 768, 1536), (92416, 768, 3072), (8192, 8192, 8192), (92416, 2048, 768), (23104, 768, 768)]:
    A = torch.randn((M, K), device="cuda").to(torch.bfloat16); Wt = torch.randn((N, K), device="cuda").to(torch.bfloat16)
    bias = torch.zeros(N, device="cuda"); out = torch.empty((M, N), dtype=torch.bfloat16, device="cuda")
    fl = 2.0 * M * N * K
    r = {}
    r["1cta"] = timeit(lambda: engine.gemm_bf16(A, Wt, bias, None, False, out=out, one_cta=True))
    r["2cta"] = timeit(lambda: engine.gemm_bf16(A, Wt, bias, None, False, out=out))
    r["m256"] = timeit(lambda: engine.gemm_bf16(A, Wt, bias, None, False, out=out, m256=True))
    Wkn = Wt.t().contiguous()
    r["cublas(A@W)"] = timeit(lambda: torch.matmul(A, Wkn, out=out))
    r["cublas(A@Wt.T)"] = timeit(lambda: torch.matmul(A, Wt.t(), out=out))
    print(f"M={M} N={N} K={K}: " + "  ".join(f"{k} {fl / v / 1e9:7.1f}" for k, v in r.items()) + "  TFLOP/s")
