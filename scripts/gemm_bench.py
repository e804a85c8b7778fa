"""Micro-benchmark of the stand-alone tcgen05 GEMM entry (kdl_gemm_bf16): TFLOP/s per shape, with cuBLAS (torch.matmul) beside it as a yardstick
(never on the product path)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import kdl_b200
from kdl_b200 import engine

def bench(M, N, K, iters=20, cublas=False):
    A = torch.randn((M, K), device="cuda").to(torch.bfloat16)
    Wt = torch.randn((N, K), device="cuda").to(torch.bfloat16)
    bias = torch.zeros(N, device="cuda")
    out = torch.empty((M, N), dtype=torch.bfloat16, device="cuda")
    for _ in range(3):
        (torch.matmul(A, Wt.t(), out=out) if cublas else engine.gemm_bf16(A, Wt, bias, None, True, out=out))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        (torch.matmul(A, Wt.t(), out=out) if cublas else engine.gemm_bf16(A, Wt, bias, None, True, out=out))
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    return ms, 2.0 * M * N * K / ms / 1e9

if __name__ == "__main__":
    shapes = [(92416, 728, 728), (92416, 768, 768), (92416, 1024, 1024), (92416, 512, 512), (65536, 2048, 2048), (92416, 728, 768), (92416, 768, 728)]
    for M, N, K in shapes:
        for name, cb in (("k_gemm_tc", False), ("cuBLAS", True)):
            ms, tf = bench(M, N, K, cublas=cb)
            print(f"M={M} N={N} K={K} {name}: {ms*1e3:8.1f} us  {tf:7.1f} TFLOP/s")
