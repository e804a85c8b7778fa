import numpy as np


def synth_inputs(h, n, seed=0):
    """uint8 U{0..255} -> the float32 tensor the gateway would send."""
    rng = np.random.default_rng(seed)
    u = rng.integers(0, 256, (n, h, h, 3), dtype=np.uint8)
    x = u.astype(np.float32)
    x /= 127.5
    x -= 1.0
    return u, x


def rel(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    d = np.max(np.abs(b))
    return float(np.max(np.abs(a - b)) / (d if d > 0 else 1.0))
