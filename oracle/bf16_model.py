"""ORACLE — test infrastructure only.  CPU model of WHERE the GPU's bf16 path rounds.

This is not reference arithmetic (the reference computes in fp32, model_server.py:55 -> TF-Serving); it restates the
storage/rounding points of csrc/engine.cu's bf16 `Runner::forward` so that the error budget of the production path can be
studied and bounded on the CPU box, where there is no GPU:

  * weights: BatchNorm folded in double, `w * gamma/sqrt(var+eps)` rounded to bf16 for conv2 / projections / pointwise
    kernels (the tcgen05 B operand); conv1, depthwise kernels, biases (folded BN shift) and the Dense head stay fp32,
  * activations: every tensor that reaches HBM is bf16 (conv outputs after bias/ReLU/residual add, depthwise outputs,
    pool+add outputs); accumulation is wider (fp32 on the GPU, fp64 here: the difference is ~1e-7, far below bf16's 4e-3),
  * the global average pool and the head run in fp32 on the bf16 block14 output.

`forward_bf16_model(W, x)` returns logits; comparing them with `xception_oracle.forward(..., float64)` gives the expected
size of the bf16 deviation (tests/test_oracle_xception.py pins it; DESIGN.md §2 quotes it).
"""
from typing import Dict, Optional

import numpy as np
import torch
import torch.nn.functional as F

from .xception_oracle import BN_EPS, _maxpool_same

D = torch.float64


def _q(t):
    """round to bf16 (round-to-nearest-even), keep computing in fp64"""
    return t.to(torch.bfloat16).to(D)


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(D)


def _fold(W, layer):
    g, b = _t(W[f"{layer}/gamma"]), _t(W[f"{layer}/beta"])
    m, v = _t(W[f"{layer}/moving_mean"]), _t(W[f"{layer}/moving_variance"])
    s = g / torch.sqrt(v + BN_EPS)
    return s, b - m * s


def _conv_folded(x, k_hwio, s, shift, stride, quant_w):
    w = _t(k_hwio) * s[None, None, None, :]
    if quant_w:
        w = _q(w.to(torch.float32).to(D))          # fp32 fold on the host, then bf16
    w = w.permute(3, 2, 0, 1).contiguous()
    return F.conv2d(x, w, stride=stride) + shift[None, :, None, None]


def _sep(x, W, layer, relu_in, relu_out, resid=None, quant_w=True, quant_t=True):
    if relu_in:
        x = F.relu(x)
    k = W[f"{layer}/depthwise_kernel"]
    t = F.conv2d(x, _t(k).permute(2, 3, 0, 1).contiguous(), padding=1, groups=k.shape[2])
    if quant_t:
        t = _q(t)                                   # depthwise output / tcgen05 A operand is bf16
    s, shift = _fold(W, f"{layer}_bn")
    y = _conv_folded(t, W[f"{layer}/pointwise_kernel"], s, shift, 1, quant_w)
    if relu_out:
        y = F.relu(y)
    if resid is not None:
        y = y + resid                               # fp32 add in the GEMM epilogue, before the rounding
    return y


def forward_bf16_model(W: Dict[str, np.ndarray], x_nhwc: np.ndarray, quant_w: bool = True, quant_act: bool = True,
                       taps: Optional[dict] = None, quant_skip: bool = True) -> np.ndarray:
    """Logits of the bf16 production path as modelled on the CPU.  quant_w / quant_act switch the two sources of rounding off
    separately (error-budget studies); quant_skip=False keeps the block outputs (the residual stream) unrounded."""
    qa = _q if quant_act else (lambda t: t)
    qx = _q if (quant_act and quant_skip) else (lambda t: t)     # block outputs = the residual (skip) stream

    def tap(name, t):
        if taps is not None:
            taps[name] = t.permute(0, 2, 3, 1).contiguous().numpy() if t.dim() == 4 else t.numpy().copy()

    with torch.no_grad():
        x = _t(x_nhwc).permute(0, 3, 1, 2).contiguous()
        s, sh = _fold(W, "block1_conv1_bn")
        x = qa(F.relu(_conv_folded(x, W["block1_conv1/kernel"], s, sh, 2, False)))    # conv1: fp32 weights, exact input
        s, sh = _fold(W, "block1_conv2_bn")
        x = qa(F.relu(_conv_folded(x, W["block1_conv2/kernel"], s, sh, 1, quant_w)))
        for b in (2, 3, 4):
            s, sh = _fold(W, f"block{b}_res_bn")
            r = qa(_conv_folded(x, W[f"block{b}_res_conv/kernel"], s, sh, 2, quant_w))
            # sepconv1 output already carries sepconv2's ReLU (folded into the producing epilogue)
            u = qa(_sep(x, W, f"block{b}_sepconv1", b != 2, True, None, quant_w, quant_act))
            tap(f"block{b}_sepconv1_bn", u)
            v = qa(_sep(u, W, f"block{b}_sepconv2", False, False, None, quant_w, quant_act))
            x = qx(_maxpool_same(v) + r)
            tap(f"block{b}_out", x)
        for b in range(5, 13):
            u = qa(_sep(x, W, f"block{b}_sepconv1", True, True, None, quant_w, quant_act))
            u = qa(_sep(u, W, f"block{b}_sepconv2", False, True, None, quant_w, quant_act))
            x = qx(_sep(u, W, f"block{b}_sepconv3", False, False, x, quant_w, quant_act))
            tap(f"block{b}_out", x)
        s, sh = _fold(W, "block13_res_bn")
        r = qa(_conv_folded(x, W["block13_res_conv/kernel"], s, sh, 2, quant_w))
        u = qa(_sep(x, W, "block13_sepconv1", True, True, None, quant_w, quant_act))
        v = qa(_sep(u, W, "block13_sepconv2", False, False, None, quant_w, quant_act))
        x = qx(_maxpool_same(v) + r)
        tap("block13_out", x)
        x = qa(_sep(x, W, "block14_sepconv1", False, True, None, quant_w, quant_act))
        x = qa(_sep(x, W, "block14_sepconv2", False, True, None, quant_w, quant_act))
        v = x.mean(dim=(2, 3))
        tap("avg_pool", v)
        if "dense_inner/kernel" in W:
            v = F.relu(v @ _t(W["dense_inner/kernel"]) + _t(W["dense_inner/bias"]))
        y = v @ _t(W["dense_7/kernel"]) + _t(W["dense_7/bias"])
        return y.numpy()
