"""Deterministic random-init weight bundles in Keras layout (no checkpoint is reachable offline:
the trained xception_v4_large_08_0.894.h5 of convert.py:4 / guide.md:176 is a download link).

Two inits:
  * ``keras_default`` -- what `keras.applications.Xception(weights=None)` + Dense layers would hold:
    glorot-uniform kernels, zero biases, identity BatchNorm.  Numerically nearly blind (middle-flow
    branches are ~1e-6 of the skip path), kept because it is "the reference's random init".
  * ``calibrated``    -- variance-preserving kernels and non-trivial BatchNorm statistics so that every
    branch is the same order as its skip path and every layer matters to the logits.

A bundle is ``dict[str, np.ndarray(float32)]`` keyed by `arch.weight_spec()` names; `pack_blob`
flattens it in canonical order for ``kdl_load_weights``.
"""
from typing import Dict

import numpy as np

from . import arch


def _fans(shape, kind):
    if kind in ("conv_kernel", "pw_kernel", "dw_kernel"):
        rf = shape[0] * shape[1]
        return rf * shape[2], rf * shape[3]
    return shape[0], shape[1]


def make_weights(seed: int = 1, init: str = "calibrated", head_inner: int = 100) -> Dict[str, np.ndarray]:
    if init not in ("keras_default", "calibrated"):
        raise ValueError(f"unknown init {init!r}")
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
    """$KDL_WEIGHTS names the weight file; without it the entry points serve the deterministic random-init bundle (the trained
    file is a download link, guide.md:176) and say so."""
    import os
    path = os.environ.get("KDL_WEIGHTS")
    if path:
        return load(path)
    print("KDL_WEIGHTS is not set: serving random-init weights (seed %d, 'calibrated')" % default_seed)
    return make_weights(default_seed, "calibrated")
