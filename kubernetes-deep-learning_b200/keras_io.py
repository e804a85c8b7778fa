"""Real-weights importer: a Keras HDF5 / .npz weight file of the clothing model -> the weight bundle
`kdl_load_weights` takes (weights.py, arch.weight_spec).

The reference's model is `keras.models.load_model('xception_v4_large_08_0.894.h5')` (convert.py:4): `keras.applications.Xception`
(include_top=False) as a nested layer, global average pooling, Dense(100, relu), [Dropout], Dense(10).  In the .h5
file every variable is a dataset `/model_weights/<layer>/<variable name>` (or `/<layer>/<variable name>` for `save_weights`), the
variable name being the TF name, e.g. `block1_conv1/kernel:0`.  Only the last two components ("<layer scope>/<variable>:0") matter
here, so whole-model files, weights-only files and nested/un-nested base models all load.

Name mapping
  * Xception's named layers (`blockN_convM`, `blockN_sepconvM`, `..._bn`) map one to one;
  * the four residual 1x1 convolutions and their BatchNorms are unnamed in keras.applications (auto names `conv2d_K`,
    `batch_normalization_K`, K depending on what was built before in the session): they are taken in increasing K, which is their
    construction order block2, block3, block4, block13 (keras/applications/xception.py);
  * Dense layers (`dense_K`) likewise in increasing K: the last one is the 10-way output (`dense_7` in the served signature),
    an optional earlier one is the inner layer; every shape is checked against `arch.weight_spec`.
The trained file itself is a download link and unreachable offline, so this path is exercised on files written here in the same
layout (`export_keras_h5`) and the HDF5 reader separately on a file written by the real HDF5 library (tests/test_hdf5_min.py);
the end-to-end golden of guide.md:623-625 remains unverified until the .h5 is available.
"""
import re
from typing import Dict, List, Tuple

import numpy as np

from . import arch, hdf5_min

_VAR = re.compile(r"^(?P<scope>[^/]+)/(?P<var>[A-Za-z_]+)(:\d+)?$")
_AUTO = re.compile(r"^(?P<base>conv2d|batch_normalization|dense)(_(?P<k>\d+))?$")
_RES_BLOCKS = (2, 3, 4, 13)
_BN_VARS = ("gamma", "beta", "moving_mean", "moving_variance")


class WeightImportError(ValueError):
    pass


def _collect(named: Dict[str, np.ndarray]) -> Dict[Tuple[str, str], np.ndarray]:
    """{'.../scope/var:0': array} -> {(scope, var): array}, ignoring optimizer state and anything that is not 'scope/var'."""
    out: Dict[Tuple[str, str], np.ndarray] = {}
    for path, a in named.items():
        parts = [p for p in path.split("/") if p]
        if "optimizer_weights" in parts or len(parts) < 2:
            continue
        m = _VAR.match("/".join(parts[-2:]))
        if not m:
            continue
        key = (m.group("scope"), m.group("var"))
        if key in out and not np.array_equal(out[key], a):
            raise WeightImportError(f"two different variables named {key[0]}/{key[1]} in the file")
        out[key] = np.asarray(a)
    return out


def bundle_from_named_arrays(named: Dict[str, np.ndarray]) -> Dict[str, np.ndarray]:
    vars_ = _collect(named)
    auto: Dict[str, List[Tuple[int, str]]] = {"conv2d": [], "batch_normalization": [], "dense": []}
    for scope in sorted({s for s, _ in vars_}):
        m = _AUTO.match(scope)
        if m:
            auto[m.group("base")].append((int(m.group("k") or 0), scope))
    for v in auto.values():
        v.sort()
    if len(auto["conv2d"]) != 4 or len(auto["batch_normalization"]) != 4:
        raise WeightImportError(f"expected the 4 unnamed residual convolutions and 4 BatchNorms of keras.applications.Xception, found "
                                f"{[s for _, s in auto['conv2d']]} and {[s for _, s in auto['batch_normalization']]}")
    if len(auto["dense"]) not in (1, 2):
        raise WeightImportError(f"expected 1 or 2 Dense layers in the head, found {[s for _, s in auto['dense']]}")
    rename: Dict[str, str] = {}
    for (_, conv), (_, bn), blk in zip(auto["conv2d"], auto["batch_normalization"], _RES_BLOCKS):
        rename[f"block{blk}_res_conv"] = conv
        rename[f"block{blk}_res_bn"] = bn
    rename["dense_7"] = auto["dense"][-1][1]
    head_inner = 0
    if len(auto["dense"]) == 2:
        rename["dense_inner"] = auto["dense"][0][1]
        k = vars_.get((rename["dense_inner"], "kernel"))
        if k is None or k.ndim != 2:
            raise WeightImportError(f"{rename['dense_inner']}/kernel is missing or not a matrix")
        head_inner = int(k.shape[1])
    out: Dict[str, np.ndarray] = {}
    used = set()
    for e in arch.weight_spec(head_inner):
        scope, var = e.name.split("/")
        src = (rename.get(scope, scope), var)
        if src not in vars_:
            raise WeightImportError(f"variable {src[0]}/{src[1]} (for {e.name}) is not in the file")
        a = vars_[src]
        if tuple(a.shape) != tuple(e.shape):
            raise WeightImportError(f"{src[0]}/{src[1]}: shape {tuple(a.shape)}, the architecture needs {tuple(e.shape)}")
        if a.dtype.kind != "f":
            raise WeightImportError(f"{src[0]}/{src[1]}: dtype {a.dtype} is not floating point")
        out[e.name] = np.ascontiguousarray(a, dtype=np.float32)
        used.add(src)
    extra = sorted(f"{s}/{v}" for (s, v) in set(vars_) - used)
    if extra:
        raise WeightImportError(f"the file holds variables the served architecture has no place for: {extra[:8]}"
                                + (" ..." if len(extra) > 8 else ""))
    return out


def load_keras_h5(path_or_bytes) -> Dict[str, np.ndarray]:
    """`xception_v4_large_08_0.894.h5` (convert.py:4) or any `model.save(...h5)` / `save_weights(...h5)` of that architecture."""
    f = hdf5_min.File(path_or_bytes)
    root = f["model_weights"] if "model_weights" in f else f
    return bundle_from_named_arrays({k: n.read() for k, n in root.walk()})


def load_keras_npz(path) -> Dict[str, np.ndarray]:
    """For a machine that has TensorFlow:  np.savez('w.npz', **{v.name: v.numpy() for v in model.variables})."""
    with np.load(path) as z:
        return bundle_from_named_arrays({k: z[k] for k in z.files})


def keras_layout(weights: Dict[str, np.ndarray], base_name: str = "xception", first_dense: int = 6,
                 first_auto: int = 4) -> Dict[str, Dict[str, np.ndarray]]:
    """bundle -> {layer: {variable name: array}} with the names Keras would give the clothing model: the base network nested as
    one layer `xception`, residual convolutions `conv2d_K` / `batch_normalization_K`, head `dense_6`, `dense_7`."""
    head_inner = weights["dense_inner/kernel"].shape[1] if "dense_inner/kernel" in weights else 0
    rename = {}
    for i, blk in enumerate(_RES_BLOCKS):
        k = first_auto + i
        rename[f"block{blk}_res_conv"] = f"conv2d_{k}" if k else "conv2d"
        rename[f"block{blk}_res_bn"] = f"batch_normalization_{k}" if k else "batch_normalization"
    d = first_dense
    if head_inner:
        rename["dense_inner"] = f"dense_{d}"
        d += 1
    rename["dense_7"] = f"dense_{d}"
    layers: Dict[str, Dict[str, np.ndarray]] = {base_name: {}}
    for e in arch.weight_spec(head_inner):
        scope, var = e.name.split("/")
        tf_scope = rename.get(scope, scope)
        layer = tf_scope if scope.startswith("dense") else base_name
        layers.setdefault(layer, {})[f"{tf_scope}/{var}:0"] = weights[e.name]
    return layers


def export_keras_h5(weights: Dict[str, np.ndarray], path=None, whole_model: bool = True) -> bytes:
    """Write the bundle as a Keras HDF5 file (`model.save` layout if whole_model, else `save_weights` layout)."""
    layers = keras_layout(weights)
    tree: dict = {".attrs": {"layer_names": np.array([n.encode() for n in layers]), "backend": np.bytes_(b"tensorflow"),
                             "keras_version": np.bytes_(b"2.4.0")}}
    for layer, vs in layers.items():
        g: dict = {".attrs": {"weight_names": np.array([n.encode() for n in vs])}}
        for name, a in vs.items():
            node = g
            parts = name.split("/")
            for p in parts[:-1]:
                node = node.setdefault(p, {})
            node[parts[-1]] = np.asarray(a, np.float32)
        tree[layer] = g
    if whole_model:
        tree = {"model_weights": tree, ".attrs": {"backend": np.bytes_(b"tensorflow"), "keras_version": np.bytes_(b"2.4.0")}}
    return hdf5_min.write_file(path, tree)
