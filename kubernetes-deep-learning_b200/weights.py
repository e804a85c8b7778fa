"""Deterministic random-init weight bundles in Keras layout (no checkpoint is reachable offline:
the trained xception_v4_large_08_0.894.h5 of convert.py:4 / guide.md:176 is a download link).

Two inits:
  * ``keras_default`` -- what `keras.applications.Xception(weights=None)` + Dense layers would hold:
    glorot-uniform kernels, zero biases, identity BatchNorm.  Numerically nearly blind (middle-flow
    branches are ~1e-6 of the skip path), kept because it is "the reference's random init".
  * ``calibrated``    -- variance-preserving kernels and non-trivial BatchNorm statistics so that every
    branch is the same order as its skip path and every layer matters to the logits.  Deliberately hostile to
    reduced precision: the residual stream doubles in variance per middle-flow block, so every branch's rounding error
    reaches the logits undamped.
  * ``calibrated_mild`` -- the same draw, with the last BatchNorm of each middle-flow branch (block5..12 sepconv3) scaled
    by 0.35, the way trained residual networks look (branches are corrections to the skip path, not its equal).  Every layer
    still matters; this is the init the literal north_star contract (bf16 logits within 2e-2 per image, identical top-1)
    is asserted on.

A bundle is ``dict[str, np.ndarray(float32)]`` keyed by `arch.weight_spec()` names; `pack_blob`
flattens it in canonical order for ``kdl_load_weights``.
"""
from typing import Dict

import numpy as np

from . import arch


MILD_GAIN = 0.35
MILD_LAYERS = tuple(f"block{b}_sepconv3_bn" for b in range(5, 13))


def _fans(shape, kind):
    if kind in ("conv_kernel", "pw_kernel", "dw_kernel"):
        rf = shape[0] * shape[1]
        return rf * shape[2], rf * shape[3]
    return shape[0], shape[1]


def make_weights(seed: int = 1, init: str = "calibrated", head_inner: int = 100) -> Dict[str, np.ndarray]:
    if init not in ("keras_default", "calibrated", "calibrated_mild"):
        raise ValueError(f"unknown init {init!r}")
    mild = init == "calibrated_mild"
    if mild:
        init = "calibrated"
    rng = np.random.default_rng(seed)
    out: Dict[str, np.ndarray] = {}
    for e in arch.weight_spec(head_inner):
        k, shape = e.kind, e.shape
        if init == "keras_default":
            if k in ("conv_kernel", "pw_kernel", "dw_kernel", "dense_kernel"):
                fi, fo = _fans(shape, k)
                lim = np.sqrt(6.0 / (fi + fo))
                a = rng.uniform(-lim, lim, size=shape)
            elif k in ("bn_gamma", "bn_var"):
                a = np.ones(shape)
            else:  # bn_beta, bn_mean, dense_bias
                a = np.zeros(shape)
        else:
            if k == "dw_kernel":
                a = rng.normal(0.0, 1.0 / 3.0, size=shape)            # var 1/9 over 9 taps
            elif k in ("conv_kernel", "pw_kernel"):
                fi, _ = _fans(shape, k)
                a = rng.normal(0.0, np.sqrt(2.0 / fi), size=shape)    # He
            elif k == "dense_kernel":
                a = rng.normal(0.0, np.sqrt(2.0 / shape[0]), size=shape)
            elif k == "bn_gamma":
                a = rng.uniform(0.75, 1.25, size=shape)
            elif k == "bn_var":
                a = rng.uniform(0.75, 1.25, size=shape)
            else:  # bn_beta, bn_mean, dense_bias
                a = rng.normal(0.0, 0.1, size=shape)
        if mild and e.kind in ("bn_gamma", "bn_beta") and e.name.split("/")[0] in MILD_LAYERS:
            a = a * MILD_GAIN
        out[e.name] = np.ascontiguousarray(a, dtype=np.float32)
    return out


def pack_blob(weights: Dict[str, np.ndarray], head_inner: int = 100) -> np.ndarray:
    """Flatten to one float32 vector in `arch.weight_spec` order (the `kdl_load_weights` blob)."""
    parts = []
    for e in arch.weight_spec(head_inner):
        a = weights[e.name]
        if tuple(a.shape) != tuple(e.shape):
            raise ValueError(f"{e.name}: shape {a.shape} != {e.shape}")
        parts.append(np.ascontiguousarray(a, dtype=np.float32).ravel())
    return np.concatenate(parts)


def unpack_blob(blob: np.ndarray, head_inner: int = 100) -> Dict[str, np.ndarray]:
    out, off = {}, 0
    for e in arch.weight_spec(head_inner):
        n = int(np.prod(e.shape))
        out[e.name] = blob[off:off + n].reshape(e.shape).copy()
        off += n
    if off != blob.size:
        raise ValueError(f"blob has {blob.size} floats, spec needs {off}")
    return out


def load(path: str) -> Dict[str, np.ndarray]:
    """Weights from a file: Keras HDF5 (`.h5`/`.hdf5`/`.keras`-era whole-model or weights-only files) or `.npz` of
    {variable name: array} (keras_io.py)."""
    from . import keras_io
    p = str(path).lower()
    if p.endswith(".npz"):
        return keras_io.load_keras_npz(path)
    if p.endswith((".h5", ".hdf5")):
        return keras_io.load_keras_h5(path)
    raise ValueError(f"unknown weight file type: {path} (expected .h5, .hdf5 or .npz)")


def load_from_env(default_seed: int = 1) -> Dict[str, np.ndarray]:
    """$KDL_WEIGHTS names the weight file the serving entry points load.  Without it they refuse to start, as the reference's
    TF-Serving container does without its model directory (tf-serving.dockerfile:5): a missing or misspelt variable must not
    give a healthy-looking service that returns meaningless logits.  KDL_ALLOW_RANDOM_WEIGHTS=1 opts into the deterministic
    random-init bundle (the trained file is a download link, guide.md:176) for demos, tests and benchmarks."""
    import os
    path = os.environ.get("KDL_WEIGHTS")
    if path:
        return load(path)
    if os.environ.get("KDL_ALLOW_RANDOM_WEIGHTS") == "1":
        print("KDL_WEIGHTS is not set: serving random-init weights (seed %d, 'calibrated') because KDL_ALLOW_RANDOM_WEIGHTS=1" % default_seed)
        return make_weights(default_seed, "calibrated")
    raise RuntimeError("KDL_WEIGHTS is not set: refusing to serve random-init weights "
                       "(point it at the Keras .h5 / .npz file, or set KDL_ALLOW_RANDOM_WEIGHTS=1 for a demo)")
