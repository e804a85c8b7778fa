"""ctypes binding of include/kdl.h.  Loading fails loudly when the CUDA library is missing: there is no
CPU fallback anywhere in this package."""
import ctypes
import os
from ctypes import POINTER, c_char_p, c_float, c_int32, c_int64, c_size_t, c_void_p

_CSRC = os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc")
LIB_PATH = os.environ.get("KDL_LIB", os.path.join(_CSRC, "libkdl.so"))   # KDL_LIB: alternative build (kernel experiments)

KDL_OK, KDL_EINVAL, KDL_ECUDA, KDL_ESTATE, KDL_ENOMEM = 0, -1, -2, -3, -4
PRECISION_F32, PRECISION_BF16 = 0, 1
GEMM_AUTO, GEMM_SIMPLE = 0, 1

# every symbol include/kdl.h declares
SYMBOLS = [
    "kdl_create", "kdl_destroy", "kdl_last_error", "kdl_weight_count", "kdl_load_weights",
    "kdl_preprocess_u8", "kdl_forward", "kdl_predict_u8", "kdl_predict_u8_host", "kdl_forward_host",
    "kdl_num_taps", "kdl_tap_name", "kdl_layer_output", "kdl_last_launch_count", "kdl_gemm_bf16",
    "kdl_version", "kdl_profile_begin", "kdl_profile_end", "kdl_preprocess_u8_static", "kdl_submit_u8_host", "kdl_wait",
    "kdl_get_tuning", "kdl_set_tuning", "kdl_graph_stats",
]
STATIC_MAX_BATCH, STATIC_MAX_SIZE = 32, 512


class KdlConfig(ctypes.Structure):
    _fields_ = [("device", c_int32), ("max_batch", c_int32), ("input_size", c_int32), ("head_inner", c_int32),
                ("precision", c_int32), ("gemm_impl", c_int32), ("keep_taps", c_int32), ("use_graphs", c_int32)]


class KdlTuning(ctypes.Structure):
    _fields_ = [("pdl_max_batch", c_int32), ("serpentine", c_int32), ("conv2_shifted_taps", c_int32), ("fused_sepconv", c_int32),
                ("fused_midflow", c_int32), ("conv1_tc", c_int32), ("dw_tma_store", c_int32), ("reserved", c_int32 * 5)]


class KdlProfileRow(ctypes.Structure):
    _fields_ = [("name", ctypes.c_char * 48), ("launches", c_int32), ("ms", c_float), ("flops", ctypes.c_double),
                ("bytes", ctypes.c_double)]


class KdlError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"kdl error {code}: {msg}")
        self.code = code


_lib = None


def load():
    """Returns the loaded library; raises if libkdl.so has not been built (run __graft_entry__.build())."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  This package has no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    P = c_void_p
    lib.kdl_create.argtypes = [POINTER(KdlConfig), POINTER(P)]
    lib.kdl_destroy.argtypes = [P]
    lib.kdl_last_error.argtypes = [P]
    lib.kdl_last_error.restype = c_char_p
    lib.kdl_weight_count.argtypes = [c_int32]
    lib.kdl_weight_count.restype = c_size_t
    lib.kdl_load_weights.argtypes = [P, P, c_size_t]
    lib.kdl_preprocess_u8.argtypes = [P, P, P, P, P, P, c_int32, P, P]
    lib.kdl_forward.argtypes = [P, P, c_int32, P, P]
    lib.kdl_predict_u8.argtypes = [P, P, P, P, P, P, c_int32, P, P]
    lib.kdl_predict_u8_host.argtypes = [P, P, c_size_t, P, P, P, P, c_int32, P]
    lib.kdl_forward_host.argtypes = [P, P, c_int32, P]
    lib.kdl_submit_u8_host.argtypes = [P, P, c_size_t, P, P, P, P, c_int32, P]
    lib.kdl_wait.argtypes = [P]
    lib.kdl_num_taps.argtypes = [P]
    lib.kdl_tap_name.argtypes = [P, c_int32]
    lib.kdl_tap_name.restype = c_char_p
    lib.kdl_layer_output.argtypes = [P, c_char_p, P, c_size_t, POINTER(c_int32 * 4), POINTER(c_int32), P]
    lib.kdl_last_launch_count.argtypes = [P]
    lib.kdl_gemm_bf16.argtypes = [P, P, P, P, P, c_int32, c_int32, c_int32, c_int32, P]
    lib.kdl_version.restype = c_char_p
    lib.kdl_preprocess_u8_static.argtypes = [P, P, P, P, P, c_int32, c_int32, P, P, c_char_p, c_size_t]
    lib.kdl_get_tuning.argtypes = [P, POINTER(KdlTuning)]
    lib.kdl_set_tuning.argtypes = [P, POINTER(KdlTuning)]
    lib.kdl_graph_stats.argtypes = [P, POINTER(c_int32), POINTER(c_int32), POINTER(c_int32), POINTER(c_int32)]
    lib.kdl_profile_begin.argtypes = [P]
    lib.kdl_profile_end.argtypes = [P, POINTER(KdlProfileRow), c_int32, POINTER(c_int32)]
    for s in SYMBOLS:
        fn = getattr(lib, s)
        if fn.restype is ctypes.c_int:
            fn.restype = c_int32
    _lib = lib
    return lib


def check(rc, handle=None):
    if rc != KDL_OK:
        msg = load().kdl_last_error(handle)
        raise KdlError(rc, msg.decode() if msg else "?")
