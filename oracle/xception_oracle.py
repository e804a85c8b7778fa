"""ORACLE — test infrastructure only (see oracle/__init__.py).  PARITY UNPINNED for the network.

torch-CPU restatement of the forward pass TF-Serving executes for model_server.py:55
(stub.Predict on model 'clothing-model' / 'serving_default', model_server.py:40-42), i.e. the graph
convert.py:4-6 exports: tf.keras.applications.Xception(include_top=False) @ tensorflow==2.3.0
(Pipfile.lock:455-471; keras/applications/xception.py) + GlobalAveragePooling2D + Dense head.
The dependency is absent from /root/reference and not installable offline, so this follows the
published architecture and TF op semantics:
  * NHWC data, cross-correlation (no kernel flip), all convs use_bias=False,
  * BatchNormalization(epsilon=1e-3) in inference mode after every conv / separable conv,
  * SeparableConv2D = depthwise 3x3 s1 SAME (zero pad 1) then pointwise 1x1, nothing in between,
  * MaxPooling2D(3, strides=2, padding='same'): TF SAME => pad_total=(out-1)*2+3-in, pad_before=pad_total//2
    (so (0,1) for even inputs, (1,1) for odd), pad value -inf,
  * residual projections Conv2D(1x1, strides=2, padding='same') sample even indices (no padding needed),
  * residual tensors are taken BEFORE the next ReLU (Keras Activation is not in place),
  * head: mean over H,W -> [Dense(inner)+bias, ReLU, Dropout=identity] -> Dense(10)+bias = raw logits
    (guide.md:622-626 shows unnormalised scores).
Run it in float64 (truth for tolerances) or float32 (stand-in for "reference TF fp32 predict").
"""
from collections import OrderedDict
from typing import Dict, Optional

import numpy as np
import torch
import torch.nn.functional as F

BN_EPS = 1e-3


def _t(a, dtype):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dtype)


def _conv(x, k_hwio, stride=1, dtype=torch.float64):
    w = _t(k_hwio, dtype).permute(3, 2, 0, 1).contiguous()
    return F.conv2d(x, w, stride=stride)


def _dw(x, k, dtype):
    c = k.shape[2]
    w = _t(k, dtype).permute(2, 3, 0, 1).contiguous()  # (C,1,3,3)
    return F.conv2d(x, w, padding=1, groups=c)


def _bn(x, W, layer, dtype):
    g = _t(W[f"{layer}/gamma"], dtype)
    b = _t(W[f"{layer}/beta"], dtype)
    m = _t(W[f"{layer}/moving_mean"], dtype)
    v = _t(W[f"{layer}/moving_variance"], dtype)
    inv = g / torch.sqrt(v + BN_EPS)
    return (x - m[None, :, None, None]) * inv[None, :, None, None] + b[None, :, None, None]


def _sep(x, W, layer, dtype):
    x = _dw(x, W[f"{layer}/depthwise_kernel"], dtype)
    x = _conv(x, W[f"{layer}/pointwise_kernel"], 1, dtype)
    return _bn(x, W, f"{layer}_bn", dtype)


def _maxpool_same(x):
    n = x.shape[-1]
    out = -(-n // 2)
    tot = max((out - 1) * 2 + 3 - n, 0)
    lo, hi = tot // 2, tot - tot // 2
    x = F.pad(x, (lo, hi, lo, hi), value=float("-inf"))
    return F.max_pool2d(x, 3, 2)


def forward(W: Dict[str, np.ndarray], x_nhwc: np.ndarray, dtype=torch.float64,
            taps: Optional[dict] = None, head_inner: Optional[int] = None) -> np.ndarray:
    """x_nhwc: (B,H,W,3) float in [-1,1] (the serving_default `input_8` tensor).  Returns (B,10) logits
    (`dense_7`).  If `taps` is a dict it is filled with NHWC numpy copies of intermediate tensors."""
    if head_inner is None:
        head_inner = W["dense_inner/kernel"].shape[1] if "dense_inner/kernel" in W else 0

    def tap(name, t):
        if taps is not None:
            taps[name] = t.permute(0, 2, 3, 1).contiguous().numpy() if t.dim() == 4 else t.numpy().copy()

    with torch.no_grad():
        x = _t(x_nhwc, dtype).permute(0, 3, 1, 2).contiguous()
        x = F.relu(_bn(_conv(x, W["block1_conv1/kernel"], 2, dtype), W, "block1_conv1_bn", dtype))
        tap("block1_conv1_act", x)
        x = F.relu(_bn(_conv(x, W["block1_conv2/kernel"], 1, dtype), W, "block1_conv2_bn", dtype))
        tap("block1_conv2_act", x)

        # entry flow blocks 2-4
        for b in (2, 3, 4):
            r = _bn(_conv(x, W[f"block{b}_res_conv/kernel"], 2, dtype), W, f"block{b}_res_bn", dtype)
            tap(f"block{b}_res_bn", r)
            if b != 2:
                x = F.relu(x)
            x = _sep(x, W, f"block{b}_sepconv1", dtype)
            tap(f"block{b}_sepconv1_bn", x)
            x = _sep(F.relu(x), W, f"block{b}_sepconv2", dtype)
            tap(f"block{b}_sepconv2_bn", x)
            x = _maxpool_same(x) + r
            tap(f"block{b}_out", x)

        # middle flow blocks 5-12
        for b in range(5, 13):
            r = x
            for j in (1, 2, 3):
                x = _sep(F.relu(x), W, f"block{b}_sepconv{j}", dtype)
                tap(f"block{b}_sepconv{j}_bn", x)
            x = x + r
            tap(f"block{b}_out", x)

        # exit flow
        r = _bn(_conv(x, W["block13_res_conv/kernel"], 2, dtype), W, "block13_res_bn", dtype)
        tap("block13_res_bn", r)
        x = _sep(F.relu(x), W, "block13_sepconv1", dtype)
        tap("block13_sepconv1_bn", x)
        x = _sep(F.relu(x), W, "block13_sepconv2", dtype)
        tap("block13_sepconv2_bn", x)
        x = _maxpool_same(x) + r
        tap("block13_out", x)
        x = F.relu(_sep(x, W, "block14_sepconv1", dtype))
        tap("block14_sepconv1_act", x)
        x = F.relu(_sep(x, W, "block14_sepconv2", dtype))
        tap("block14_sepconv2_act", x)

        v = x.mean(dim=(2, 3))
        tap("avg_pool", v)
        if head_inner > 0:
            v = F.relu(v @ _t(W["dense_inner/kernel"], dtype) + _t(W["dense_inner/bias"], dtype))
            tap("dense_inner", v)
        y = v @ _t(W["dense_7/kernel"], dtype) + _t(W["dense_7/bias"], dtype)
        tap("dense_7", y)
        return y.numpy()


def rel_err(a: np.ndarray, b: np.ndarray) -> float:
    """max|a-b| / max|b| over the whole vector/tensor (never element-wise)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    d = float(np.max(np.abs(b)))
    return float(np.max(np.abs(a - b))) / (d if d > 0 else 1.0)
