import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, kdl_b200
from kdl_b200 import engine
M, N, K = 92416, 768, 768
A = torch.randn((M, K), device="cuda").to(torch.bfloat16); Wt = torch.randn((N, K), device="cuda").to(torch.bfloat16)
bias = torch.zeros(N, device="cuda"); out = torch.empty((M, N), dtype=torch.bfloat16, device="cuda")
for _ in range(6):
    engine.gemm_bf16(A, Wt, bias, None, True, out=out)
torch.cuda.synchronize()
