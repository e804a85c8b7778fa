"""Architecture spec of the served model: Keras Xception (include_top=False) + the clothing head.

This is the graph convert.py:4-6 exports and TF-Serving executes behind model_server.py:55
(signature guide.md:220-231: input_8 (-1,299,299,3) f32 -> dense_7 (-1,10) f32).  The reference
holds none of the graph's arithmetic in-tree (it lives in tensorflow==2.3.0,
keras/applications/xception.py); the layer list below restates that published architecture.

Only shapes / names / ordering live here -- no arithmetic.  The CUDA engine (csrc/) builds the same
table in C++; `weight_spec()` is the canonical weight order shared by `weights.pack_blob` and
`kdl_load_weights`.
"""
from dataclasses import dataclass
from typing import List, Tuple

BN_EPS = 1e-3  # Keras BatchNormalization(epsilon=1e-3) as used by applications.Xception

# (block, [(cin, cout) for each sepconv], has_residual_projection, pool_at_end)
ENTRY_EXIT_BLOCKS = {
    2: ([(64, 128), (128, 128)], True),
    3: ([(128, 256), (256, 256)], True),
    4: ([(256, 728), (728, 728)], True),
    13: ([(728, 728), (728, 1024)], True),
}
MIDDLE_BLOCKS = list(range(5, 13))  # 8 blocks of 3 x sepconv 728->728, identity residual
EXIT_SEPCONVS = [(1024, 1536), (1536, 2048)]  # block14, ReLU *after* BN


def conv_out_valid(n: int, k: int, s: int) -> int:
    return (n - k) // s + 1


def same_out(n: int, s: int) -> int:
    return -(-n // s)


def spatial_chain(h: int) -> List[int]:
    """299 -> [299,149,147,74,37,19,10];  150 -> [150,74,72,36,18,9,5]."""
    a = conv_out_valid(h, 3, 2)
    b = conv_out_valid(a, 3, 1)
    c = same_out(b, 2)
    d = same_out(c, 2)
    e = same_out(d, 2)
    f = same_out(e, 2)
    return [h, a, b, c, d, e, f]


@dataclass(frozen=True)
class WeightEntry:
    name: str          # Keras-style "<layer>/<variable>"
    shape: Tuple[int, ...]
    kind: str          # conv_kernel | dw_kernel | pw_kernel | bn_gamma | bn_beta | bn_mean | bn_var | dense_kernel | dense_bias


def _bn(layer: str, c: int) -> List[WeightEntry]:
    return [
        WeightEntry(f"{layer}/gamma", (c,), "bn_gamma"),
        WeightEntry(f"{layer}/beta", (c,), "bn_beta"),
        WeightEntry(f"{layer}/moving_mean", (c,), "bn_mean"),
        WeightEntry(f"{layer}/moving_variance", (c,), "bn_var"),
    ]


def _sep(layer: str, cin: int, cout: int) -> List[WeightEntry]:
    return [
        WeightEntry(f"{layer}/depthwise_kernel", (3, 3, cin, 1), "dw_kernel"),
        WeightEntry(f"{layer}/pointwise_kernel", (1, 1, cin, cout), "pw_kernel"),
    ] + _bn(f"{layer}_bn", cout)


def weight_spec(head_inner: int = 100, n_classes: int = 10) -> List[WeightEntry]:
    """Canonical (blob) order.  Layouts are Keras': conv HWIO, depthwise (3,3,C,1), dense (in,out)."""
    w: List[WeightEntry] = []
    w += [WeightEntry("block1_conv1/kernel", (3, 3, 3, 32), "conv_kernel")] + _bn("block1_conv1_bn", 32)
    w += [WeightEntry("block1_conv2/kernel", (3, 3, 32, 64), "conv_kernel")] + _bn("block1_conv2_bn", 64)
    for b in (2, 3, 4):
        seps, _ = ENTRY_EXIT_BLOCKS[b]
        cin, cout = seps[0][0], seps[-1][1]
        w += [WeightEntry(f"block{b}_res_conv/kernel", (1, 1, cin, cout), "conv_kernel")] + _bn(f"block{b}_res_bn", cout)
        for j, (ci, co) in enumerate(seps, 1):
            w += _sep(f"block{b}_sepconv{j}", ci, co)
    for b in MIDDLE_BLOCKS:
        for j in (1, 2, 3):
            w += _sep(f"block{b}_sepconv{j}", 728, 728)
    seps, _ = ENTRY_EXIT_BLOCKS[13]
    w += [WeightEntry("block13_res_conv/kernel", (1, 1, 728, 1024), "conv_kernel")] + _bn("block13_res_bn", 1024)
    for j, (ci, co) in enumerate(seps, 1):
        w += _sep(f"block13_sepconv{j}", ci, co)
    for j, (ci, co) in enumerate(EXIT_SEPCONVS, 1):
        w += _sep(f"block14_sepconv{j}", ci, co)
    feat = 2048
    if head_inner > 0:
        w += [WeightEntry("dense_inner/kernel", (feat, head_inner), "dense_kernel"),
              WeightEntry("dense_inner/bias", (head_inner,), "dense_bias")]
        feat = head_inner
    w += [WeightEntry("dense_7/kernel", (feat, n_classes), "dense_kernel"),
          WeightEntry("dense_7/bias", (n_classes,), "dense_bias")]
    return w


def param_count(head_inner: int = 100, base_only: bool = False) -> int:
    n = 0
    for e in weight_spec(head_inner):
        if base_only and e.name.startswith("dense"):
            continue
        k = 1
        for d in e.shape:
            k *= d
        n += k
    return n


def mac_count(h: int = 299, head_inner: int = 100) -> int:
    """MACs per image (8 355 561 296 @299, 1 924 603 256 @150)."""
    s = spatial_chain(h)
    macs = s[1] ** 2 * 27 * 32 + s[2] ** 2 * 288 * 64

    def sep(hw, ci, co):
        return hw * hw * (9 * ci + ci * co)

    macs += s[3] ** 2 * 64 * 128 + sep(s[2], 64, 128) + sep(s[2], 128, 128)
    macs += s[4] ** 2 * 128 * 256 + sep(s[3], 128, 256) + sep(s[3], 256, 256)
    macs += s[5] ** 2 * 256 * 728 + sep(s[4], 256, 728) + sep(s[4], 728, 728)
    macs += 24 * sep(s[5], 728, 728)
    macs += s[6] ** 2 * 728 * 1024 + sep(s[5], 728, 728) + sep(s[5], 728, 1024)
    macs += sep(s[6], 1024, 1536) + sep(s[6], 1536, 2048)
    macs += 2048 * head_inner + head_inner * 10 if head_inner > 0 else 2048 * 10
    return macs
