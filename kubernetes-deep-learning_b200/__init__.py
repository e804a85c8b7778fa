"""B200-native drop-in for the one hot path of alexeygrigorev/kubernetes-deep-learning:
gateway preprocessing (model_server.py:18,53) + the clothing-model Xception-299 forward that
TF-Serving runs behind model_server.py:55.  Host code is Python calling a C-ABI CUDA library
(include/kdl.h, csrc/); there is no CPU fallback."""
from . import arch, weights  # noqa: F401

__all__ = ["arch", "weights", "engine", "gateway", "replicas", "serving", "http_gateway", "batching", "hdf5_min", "keras_io"]


def __getattr__(name):
    # engine / gateway / replicas import torch + the CUDA library: load lazily so that arch/weights stay light
    if name in ("engine", "gateway", "replicas", "build", "_lib", "serving", "http_gateway", "batching", "hdf5_min", "keras_io"):
        import importlib
        return importlib.import_module(f"{__name__}.{name}")
    raise AttributeError(name)
