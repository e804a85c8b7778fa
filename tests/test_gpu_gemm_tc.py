"""Kernel-level parity of the tcgen05/TMEM/TMA pointwise GEMM (kdl_gemm_bf16) against a torch fp32 reference
of the same op on the same bf16 inputs.  Tolerance: bf16 output rounding (2^-8 relative) + fp32 accumulation
order -> |got - ref| <= 1e-2 * max|ref|."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

SHAPES = [
    # (M, N, K) : the network's pointwise / projection shapes at small batch + ragged edge cases
    (361, 728, 728), (2 * 361, 728, 728), (1369, 728, 256), (100, 1024, 728), (100, 1536, 1024), (200, 2048, 1536),
    (21609, 128, 64), (5476, 256, 128), (128, 256, 64), (1, 8, 8), (127, 264, 72), (129, 248, 200), (4096, 728, 728),
    (256, 728, 728), (257, 40, 64), (92416, 728, 728), (1000, 1000, 1000),
]


@pytest.mark.parametrize("M,N,K", SHAPES)
@pytest.mark.parametrize("relu,resid", [(False, False), (True, False), (False, True)])
def test_gemm_tc_matches_fp32_reference(M, N, K, relu, resid):
    import kdl_b200  # noqa: F401
    from kdl_b200 import engine
    g = torch.Generator(device="cuda").manual_seed(M * 31 + N * 7 + K)
    A = torch.randn((M, K), generator=g, device="cuda").to(torch.bfloat16)
    Wt = (torch.randn((N, K), generator=g, device="cuda") / K ** 0.5).to(torch.bfloat16)
    bias = torch.randn((N,), generator=g, device="cuda")
    R = torch.randn((M, N), generator=g, device="cuda").to(torch.bfloat16) if resid else None
    out = engine.gemm_bf16(A, Wt, bias, R, relu)
    torch.cuda.synchronize()
    ref = A.float() @ Wt.float().t() + bias
    if relu:
        ref = torch.relu(ref)
    if resid:
        ref = ref + R.float()
    err = (out.float() - ref).abs().max().item() / ref.abs().max().item()
    assert err <= 1e-2, err


def test_gemm_tc_in_place_residual():
    import kdl_b200  # noqa: F401
    from kdl_b200 import engine
    g = torch.Generator(device="cuda").manual_seed(5)
    M, N, K = 722, 728, 728
    A = torch.randn((M, K), generator=g, device="cuda").to(torch.bfloat16)
    Wt = (torch.randn((N, K), generator=g, device="cuda") / K ** 0.5).to(torch.bfloat16)
    bias = torch.randn((N,), generator=g, device="cuda")
    X = torch.randn((M, N), generator=g, device="cuda").to(torch.bfloat16)
    ref = A.float() @ Wt.float().t() + bias + X.float()
    engine.gemm_bf16(A, Wt, bias, X, False, out=X)     # the middle-flow skip connection is updated in place
    torch.cuda.synchronize()
    assert (X.float() - ref).abs().max().item() / ref.abs().max().item() <= 1e-2
