"""ORACLE — test infrastructure only (see oracle/__init__.py).

numpy restatement of the gateway preprocessing, plus loaders for the plain-C restatement
(oracle/preprocess_oracle.c) and for the reference's real implementation (Pillow + numpy,
exactly the calls keras-image-helper 0.0.1 makes behind model_server.py:18,53).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libkdl_oracle.so")


def index_table(n_in: int, n_out: int) -> np.ndarray:
    """Pillow 8.1.0 libImaging/Geometry.c ImagingScaleAffine index rule (reached from
    model_server.py:53 -> keras-image-helper resize_image -> Image.resize(size, NEAREST)):
    a = n_in/n_out in double, xo = a*0.5, idx = (int)xo, xo += a  (accumulated)."""
    a = float(n_in) / float(n_out)
    o = a * 0.5
    tab = np.empty(n_out, dtype=np.int32)
    for i in range(n_out):
        tab[i] = -1 if o < 0.0 else int(o)
        o += a
    return tab


def resize_nearest_np(img: np.ndarray, out_h: int = 299, out_w: int = 299) -> np.ndarray:
    """uint8 HWC -> uint8 (out_h, out_w, 3) gather with the tables above."""
    assert img.dtype == np.uint8 and img.ndim == 3 and img.shape[2] == 3
    yt = index_table(img.shape[0], out_h)
    xt = index_table(img.shape[1], out_w)
    return np.ascontiguousarray(img[yt][:, xt])


def scale_np(u8: np.ndarray) -> np.ndarray:
    """keras-image-helper 0.0.1 convert_to_tensor + tf-mode preprocessing (model_server.py:18):
    np.array(img, dtype='float32'); expand_dims; x /= 127.5; x -= 1.0 (in place, fp32)."""
    x = np.array(u8, dtype="float32")
    x = np.expand_dims(x, axis=0)
    x /= 127.5
    x -= 1.0
    return x


def preprocess_np(img: np.ndarray, out_h: int = 299, out_w: int = 299) -> np.ndarray:
    return scale_np(resize_nearest_np(img, out_h, out_w))


def preprocess_pillow(img: np.ndarray, out_h: int = 299, out_w: int = 299) -> np.ndarray:
    """The reference's real code path: PIL Image.resize((w,h), NEAREST) + numpy fp32 scaling.
    This is the `"kind": "reference"` CPU baseline for preprocessing."""
    from PIL import Image

    pil = Image.fromarray(img)
    if pil.mode != "RGB":
        pil = pil.convert("RGB")
    pil = pil.resize((out_w, out_h), Image.NEAREST)
    return scale_np(np.asarray(pil))


def resize_pillow(img: np.ndarray, out_h: int = 299, out_w: int = 299) -> np.ndarray:
    from PIL import Image

    return np.asarray(Image.fromarray(img).resize((out_w, out_h), Image.NEAREST))


# ---------------------------------------------------------------------------------------------
# plain-C restatement
# ---------------------------------------------------------------------------------------------
def build_c(force: bool = False) -> str:
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(
        os.path.join(_HERE, "preprocess_oracle.c")
    ):
        subprocess.check_call(["make", "-s", "-C", _HERE])
    return _SO


_lib = None


def c_lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build_c()
        lib = ctypes.CDLL(_SO)
        lib.kdl_oracle_index_table.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
        lib.kdl_oracle_index_table.restype = ctypes.c_int
        lib.kdl_oracle_preprocess.argtypes = [
            ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int64,
            ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
            ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
        ]
        lib.kdl_oracle_preprocess.restype = ctypes.c_int
        lib.kdl_oracle_resize_u8.argtypes = [
            ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int64,
            ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
        ]
        lib.kdl_oracle_resize_u8.restype = ctypes.c_int
        _lib = lib
    return _lib


def index_table_c(n_in: int, n_out: int) -> np.ndarray:
    tab = np.empty(n_out, dtype=np.int32)
    rc = c_lib().kdl_oracle_index_table(n_in, n_out, tab.ctypes.data)
    assert rc == 0
    return tab


def resize_c(img: np.ndarray, out_h: int = 299, out_w: int = 299) -> np.ndarray:
    img = np.ascontiguousarray(img)
    h, w, _ = img.shape
    out = np.empty((out_h, out_w, 3), dtype=np.uint8)
    xt = np.empty(out_w, dtype=np.int32)
    yt = np.empty(out_h, dtype=np.int32)
    rc = c_lib().kdl_oracle_resize_u8(img.ctypes.data, h, w, w * 3, out.ctypes.data, out_h, out_w,
                                      xt.ctypes.data, yt.ctypes.data)
    assert rc == 0
    return out


def preprocess_c(img: np.ndarray, out_h: int = 299, out_w: int = 299) -> np.ndarray:
    img = np.ascontiguousarray(img)
    h, w, _ = img.shape
    out = np.empty((1, out_h, out_w, 3), dtype=np.float32)
    u8 = np.empty((out_h, out_w, 3), dtype=np.uint8)
    xt = np.empty(out_w, dtype=np.int32)
    yt = np.empty(out_h, dtype=np.int32)
    rc = c_lib().kdl_oracle_preprocess(img.ctypes.data, h, w, w * 3, out.ctypes.data, out_h, out_w,
                                       u8.ctypes.data, xt.ctypes.data, yt.ctypes.data)
    assert rc == 0
    return out
