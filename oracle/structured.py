"""ORACLE-side test inputs — test infrastructure only (see oracle/__init__.py).

Why this exists: i.i.d. U{0..255} noise images are statistically identical, so the
logits of a random-init network barely depend on which noise image it saw and "identical top-1" can be
passed by a kernel that ignores the image.  The parity tests therefore use pictures with structure
(the reference's own media/image2.png and crops/flips of it, solid colours, ramps, checkerboards,
low-frequency fields) and a head whose biases are re-centred on that batch, so that the logits are
dominated by the image-dependent part and the batch covers several argmax classes.

Nothing here is reference arithmetic; model_server.py:53 is the call whose inputs these imitate
(the gateway sees decoded RGB pictures of arbitrary size).
"""
import os

import numpy as np

_GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _lowfreq(rng, h, w, k):
    """Smooth random colour field: a k x k x 3 Gaussian grid, bilinearly interpolated to h x w (numpy only)."""
    g = rng.normal(size=(k, k, 3))
    ys = np.linspace(0, k - 1, h)
    xs = np.linspace(0, k - 1, w)
    y0 = np.clip(np.floor(ys).astype(int), 0, k - 2)
    x0 = np.clip(np.floor(xs).astype(int), 0, k - 2)
    fy = (ys - y0)[:, None, None]
    fx = (xs - x0)[None, :, None]
    a = (g[y0][:, x0] * (1 - fy) * (1 - fx) + g[y0][:, x0 + 1] * (1 - fy) * fx +
         g[y0 + 1][:, x0] * fy * (1 - fx) + g[y0 + 1][:, x0 + 1] * fy * fx)
    a = (a - a.min()) / (a.max() - a.min())
    return (a * 255.0).astype(np.uint8)


def _checker(h, w, period):
    yy, xx = np.indices((h, w))
    v = (((yy // period) + (xx // period)) % 2 * 255).astype(np.uint8)
    return np.repeat(v[:, :, None], 3, 2)


def _ramp(h, w, axis, rgb_gain=(1.0, 1.0, 1.0)):
    n = w if axis == 1 else h
    r = np.linspace(0, 255, n)
    r = r[None, :] if axis == 1 else r[:, None]
    a = np.broadcast_to(r, (h, w))[:, :, None] * np.asarray(rgb_gain)[None, None, :]
    return a.astype(np.uint8)


def structured_images(n, seed=0, mixed_sizes=True, size=299):
    """n uint8 HWC pictures, deterministic in (n, seed).  With mixed_sizes the sources have assorted sizes (the resize runs for
    real); otherwise every picture is size x size.  The first 16 are fixed kinds; further ones are low-frequency fields, crops of
    the fixture and colour-tinted checkerboards drawn from `seed`."""
    rng = np.random.default_rng(seed)
    src = np.load(os.path.join(_GOLDEN, "image2_fixture.npz"))["src"]     # media/image2.png decoded (534 x 400 x 3)

    def dims(h, w):
        return (h, w) if mixed_sizes else (size, size)

    def fit(img):
        if mixed_sizes:
            return np.ascontiguousarray(img)
        yi = (np.arange(size) * img.shape[0] // size)
        xi = (np.arange(size) * img.shape[1] // size)
        return np.ascontiguousarray(img[yi][:, xi])

    fixed = [
        lambda: fit(src),
        lambda: fit(src[::-1]),
        lambda: fit(src[:, ::-1]),
        lambda: fit(src[100:400, 50:350]),
        lambda: np.full(dims(299, 299) + (3,), 0, np.uint8),
        lambda: np.full(dims(240, 320) + (3,), 255, np.uint8),
        lambda: np.full(dims(300, 200) + (3,), (200, 30, 30), np.uint8),
        lambda: np.full(dims(64, 64) + (3,), (20, 40, 220), np.uint8),
        lambda: _ramp(*dims(300, 400), 1),
        lambda: _ramp(*dims(500, 320), 0, (1.0, 0.5, 0.1)),
        lambda: _checker(*dims(299, 299), 8),
        lambda: _checker(*dims(600, 450), 32),
        lambda: _lowfreq(rng, *dims(299, 299), 6),
        lambda: _lowfreq(rng, *dims(480, 640), 4),
        lambda: _lowfreq(rng, *dims(200, 800), 10),
        lambda: fit(255 - src),
    ]
    out = []
    for i in range(n):
        if i < len(fixed):
            out.append(fixed[i]())
            continue
        kind = i % 4
        h, w = (int(rng.integers(64, 700)), int(rng.integers(64, 700)))
        if kind == 0:
            out.append(_lowfreq(rng, *dims(h, w), int(rng.integers(3, 14))))
        elif kind == 1:
            y0, x0 = int(rng.integers(0, 300)), int(rng.integers(0, 200))
            c = src[y0:y0 + int(rng.integers(120, 234)), x0:x0 + int(rng.integers(100, 200))]
            if rng.integers(0, 2):
                c = c[:, ::-1]
            out.append(fit(c))
        elif kind == 2:
            tint = rng.uniform(0.1, 1.0, 3)
            c = _checker(*dims(h, w), int(rng.integers(2, 64))).astype(np.float64) * tint
            out.append(c.astype(np.uint8))
        else:
            base = _lowfreq(rng, *dims(h, w), int(rng.integers(2, 6))).astype(np.int32)
            noise = rng.integers(-24, 25, base.shape)
            out.append(np.clip(base + noise, 0, 255).astype(np.uint8))
    return out


def recentre_head(W, avg_pool):
    """Returns a copy of the weight bundle whose two head biases are re-centred on this batch of pooled features
    (avg_pool: (B, 2048) float64 from the fp64 oracle): dense_inner pre-activations get zero mean over the batch (about half
    the inner units fire for any picture) and the logits get zero mean per class.  Only the biases change, so the bundle
    is still a legal Keras weight set; the image-independent common mode that swamped the logits of the plain 'calibrated'
    init is gone."""
    W2 = dict(W)
    v = np.asarray(avg_pool, np.float64)
    if "dense_inner/kernel" in W:
        pre = v @ W["dense_inner/kernel"].astype(np.float64)
        b1 = -pre.mean(0)
        W2["dense_inner/bias"] = b1.astype(np.float32)
        hid = np.maximum(pre + W2["dense_inner/bias"].astype(np.float64), 0.0)
    else:
        hid = v
    y = hid @ W["dense_7/kernel"].astype(np.float64)
    W2["dense_7/bias"] = (-y.mean(0)).astype(np.float32)
    return W2


def head_only(W, avg_pool):
    """fp64 logits from pooled features (the oracle's head, xception_oracle.forward's last lines) — lets a test re-evaluate the
    head after recentre_head without a second pass through the network."""
    v = np.asarray(avg_pool, np.float64)
    if "dense_inner/kernel" in W:
        v = np.maximum(v @ W["dense_inner/kernel"].astype(np.float64) + W["dense_inner/bias"].astype(np.float64), 0.0)
    return v @ W["dense_7/kernel"].astype(np.float64) + W["dense_7/bias"].astype(np.float64)


def rel(a, b):
    """north_star's per-image metric: max|a-b| / max|b| over one logit vector."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    d = float(np.max(np.abs(b)))
    return float(np.max(np.abs(a - b))) / (d if d > 0 else 1.0)


def centred_rel(a, b):
    """Per-image error of the image-DEPENDENT part: subtract the batch mean (per class) from both sides first, so that a large
    common mode in the logits cannot hide an error in the part that distinguishes pictures."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    ac = a - a.mean(0, keepdims=True)
    bc = b - b.mean(0, keepdims=True)
    return [rel(ac[i], bc[i]) for i in range(len(a))]
