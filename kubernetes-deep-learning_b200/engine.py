"""Host-side model object over the C ABI.

`Model` is one GPU replica of the served graph.  Its three entry points mirror the reference's seams:
  * `serving_default(input_8=...) -> {'dense_7': ...}`  -- the SavedModel signature TF-Serving exposes
    (guide.md:220-231; called through model_server.py:38-43,55),
  * `predict_uint8(images)`                              -- apply_model without the RPC hop
    (model_server.py:52-56): decoded RGB uint8 images of any size -> (B,10) logits,
  * `preprocess_uint8(images)`                           -- preprocessor.from_url minus download/decode
    (model_server.py:18,53): -> float32 (B,S,S,3), bit-exact with Pillow NEAREST + x/127.5-1.

torch is used for device memory and streams only.  No CPU fallback: constructing a Model without a CUDA
device or without the built library raises.
"""
import ctypes
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from . import _lib, arch
from .weights import pack_blob

MODEL_NAME = "clothing-model"          # model_server.py:40, tf-serving.dockerfile:4
SIGNATURE_NAME = "serving_default"     # model_server.py:41
INPUT_KEY = "input_8"                  # model_server.py:42, guide.md:222
OUTPUT_KEY = "dense_7"                 # model_server.py:47, guide.md:227


class PackedImages:
    """B decoded RGB uint8 images of arbitrary sizes packed into one (pinned) host buffer, plus the
    per-image metadata arrays the C ABI takes.  Offsets are 256-byte aligned."""

    def __init__(self, images: Sequence[np.ndarray], pinned: bool = True, staging: Optional[torch.Tensor] = None):
        """`staging`: an existing (pinned) uint8 CPU tensor to pack into instead of allocating one -- pinning memory costs far more
        than a batch-1 forward pass, so request paths reuse one buffer (Model.pack)."""
        self.B = len(images)
        self.h = np.empty(self.B, np.int32)
        self.w = np.empty(self.B, np.int32)
        self.pitch = np.empty(self.B, np.int64)
        self.offset = np.empty(self.B, np.int64)
        off = 0
        for i, im in enumerate(images):
            if im.dtype != np.uint8 or im.ndim != 3 or im.shape[2] != 3:
                raise ValueError(f"image {i}: expected uint8 HWC RGB, got {im.dtype} {im.shape}")
            self.h[i], self.w[i] = im.shape[0], im.shape[1]
            self.pitch[i] = im.shape[1] * 3
            self.offset[i] = off
            off += (im.shape[0] * im.shape[1] * 3 + 255) // 256 * 256
        self.total_bytes = max(off, 256)
        if staging is not None and staging.numel() >= self.total_bytes:
            self.host = staging
        else:
            self.host = torch.empty(self.total_bytes, dtype=torch.uint8, pin_memory=pinned and torch.cuda.is_available())
        hv = self.host.numpy()
        for i, im in enumerate(images):
            n = im.shape[0] * im.shape[1] * 3
            hv[self.offset[i]:self.offset[i] + n] = np.ascontiguousarray(im).reshape(-1)

    def meta_ptrs(self):
        return (self.offset.ctypes.data, self.h.ctypes.data, self.w.ctypes.data, self.pitch.ctypes.data)


class Model:
    def __init__(self, weights: Dict[str, np.ndarray], device: int = 0, max_batch: int = 64, input_size: int = 299,
                 precision: str = "bf16", gemm_impl: str = "auto", keep_taps: bool = False,
                 head_inner: Optional[int] = None, use_graphs: bool = True, tuning: Optional[Dict[str, int]] = None):
        if not torch.cuda.is_available():
            raise RuntimeError("kdl_b200.Model needs a CUDA device (B200, sm_100a); there is no CPU fallback")
        self.lib = _lib.load()
        if head_inner is None:
            head_inner = weights["dense_inner/kernel"].shape[1] if "dense_inner/kernel" in weights else 0
        self.device = device
        self.max_batch = max_batch
        self.input_size = input_size
        self.precision = precision
        self.head_inner = head_inner
        cfg = _lib.KdlConfig(device=device, max_batch=max_batch, input_size=input_size, head_inner=head_inner,
                             precision={"f32": _lib.PRECISION_F32, "bf16": _lib.PRECISION_BF16}[precision],
                             gemm_impl={"auto": _lib.GEMM_AUTO, "simple": _lib.GEMM_SIMPLE}[gemm_impl],
                             keep_taps=int(keep_taps), use_graphs=int(use_graphs and not keep_taps))
        h = ctypes.c_void_p()
        torch.cuda.init()
        with torch.cuda.device(device):
            _lib.check(self.lib.kdl_create(ctypes.byref(cfg), ctypes.byref(h)))
            self.h = h
            blob = pack_blob(weights, head_inner)
            assert blob.size == self.lib.kdl_weight_count(head_inner)
            _lib.check(self.lib.kdl_load_weights(self.h, blob.ctypes.data, blob.size), self.h)
        if tuning:
            self.set_tuning(**tuning)

    def get_tuning(self) -> Dict[str, int]:
        t = _lib.KdlTuning()
        _lib.check(self.lib.kdl_get_tuning(self.h, ctypes.byref(t)), self.h)
        return {name: int(getattr(t, name)) for name, _ in _lib.KdlTuning._fields_ if name != "reserved"}

    def set_tuning(self, **switches: int) -> None:
        """Kernel-selection switches of include/kdl.h KdlTuning (e.g. fused_sepconv=0, pdl_max_batch=0); results are bit-identical
        whatever is selected."""
        t = _lib.KdlTuning()
        _lib.check(self.lib.kdl_get_tuning(self.h, ctypes.byref(t)), self.h)
        for k, v in switches.items():
            if k == "reserved" or not hasattr(t, k):
                raise KeyError(f"unknown tuning switch {k!r}")
            setattr(t, k, int(v))
        _lib.check(self.lib.kdl_set_tuning(self.h, ctypes.byref(t)), self.h)

    def graph_stats(self) -> Dict[str, int]:
        v = [ctypes.c_int32() for _ in range(4)]
        _lib.check(self.lib.kdl_graph_stats(self.h, *[ctypes.byref(x) for x in v]), self.h)
        return dict(zip(("cached", "captures", "replays", "evictions"), (int(x.value) for x in v)))

    def close(self):
        if getattr(self, "h", None):
            self.lib.kdl_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------------------------------
    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _dev(self):
        return torch.device("cuda", self.device)

    @property
    def last_launch_count(self) -> int:
        return int(self.lib.kdl_last_launch_count(self.h))

    def forward_device(self, x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """x: float32 CUDA tensor (B,S,S,3) in [-1,1]  ->  float32 CUDA tensor (B,10).  Async on the current stream."""
        S = self.input_size
        if x.dtype != torch.float32 or not x.is_cuda or x.dim() != 4 or tuple(x.shape[1:]) != (S, S, 3):
            raise ValueError(f"input_8 must be a float32 CUDA tensor (-1,{S},{S},3), got {x.dtype} {tuple(x.shape)}")
        x = x.contiguous()
        B = x.shape[0]
        if out is None:
            out = torch.empty((B, 10), dtype=torch.float32, device=x.device)
        with torch.cuda.device(self.device):
            _lib.check(self.lib.kdl_forward(self.h, x.data_ptr(), B, out.data_ptr(), self._stream()), self.h)
        return out

    def serving_default(self, **inputs) -> Dict[str, np.ndarray]:
        """The SavedModel signature (guide.md:220-231): inputs['input_8'] float32 (-1,299,299,3) ->
        outputs['dense_7'] float32 (-1,10).  Host arrays in, host arrays out."""
        if set(inputs) != {INPUT_KEY}:
            raise KeyError(f"signature '{SIGNATURE_NAME}' takes exactly one input named '{INPUT_KEY}'")
        x = np.ascontiguousarray(inputs[INPUT_KEY], dtype=np.float32)
        S = self.input_size
        if x.ndim != 4 or x.shape[1:] != (S, S, 3):
            raise ValueError(f"{INPUT_KEY} must have shape (-1,{S},{S},3), got {x.shape}")
        outs = []
        for i in range(0, x.shape[0], self.max_batch):
            xb = x[i:i + self.max_batch]
            y = np.empty((xb.shape[0], 10), np.float32)
            _lib.check(self.lib.kdl_forward_host(self.h, xb.ctypes.data, xb.shape[0], y.ctypes.data), self.h)
            outs.append(y)
        return {OUTPUT_KEY: np.concatenate(outs, 0) if outs else np.empty((0, 10), np.float32)}

    def preprocess_uint8(self, images: Sequence[np.ndarray]) -> torch.Tensor:
        """-> float32 CUDA tensor (B,S,S,3): Pillow-NEAREST resize + x/127.5-1, bit-exact."""
        pk = PackedImages(images)
        d_src = pk.host.to(self._dev(), non_blocking=True)
        S = self.input_size
        out = torch.empty((pk.B, S, S, 3), dtype=torch.float32, device=self._dev())
        with torch.cuda.device(self.device):
            _lib.check(self.lib.kdl_preprocess_u8(self.h, d_src.data_ptr(), *pk.meta_ptrs(), pk.B, out.data_ptr(),
                                                  self._stream()), self.h)
        torch.cuda.current_stream(self.device).synchronize()
        return out

    def predict_packed_device(self, d_src: torch.Tensor, pk: PackedImages, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Source pixels already resident on the device (d_src = pk.host copied).  Async."""
        if out is None:
            out = torch.empty((pk.B, 10), dtype=torch.float32, device=self._dev())
        with torch.cuda.device(self.device):
            _lib.check(self.lib.kdl_predict_u8(self.h, d_src.data_ptr(), *pk.meta_ptrs(), pk.B, out.data_ptr(),
                                               self._stream()), self.h)
        return out

    def predict_packed_host(self, pk: PackedImages, out: Optional[np.ndarray] = None) -> np.ndarray:
        """End to end from (pinned) host memory: H2D of the packed pixels, fused resize+scale, forward,
        D2H of the logits; synchronous."""
        if out is None:
            out = np.empty((pk.B, 10), np.float32)
        _lib.check(self.lib.kdl_predict_u8_host(self.h, pk.host.data_ptr(), pk.total_bytes, *pk.meta_ptrs(), pk.B,
                                                out.ctypes.data), self.h)
        return out

    def submit_packed_host(self, pk: PackedImages, out: torch.Tensor) -> None:
        """Pipelined end-to-end call: queues H2D + kernels + D2H and returns; `out` is a pinned float32 CPU tensor (B,10) that
        is valid after the matching `wait()`.  At most two submits are in flight (the copy of one overlaps the kernels of the other)."""
        assert out.dtype == torch.float32 and not out.is_cuda and out.numel() >= pk.B * 10
        _lib.check(self.lib.kdl_submit_u8_host(self.h, pk.host.data_ptr(), pk.total_bytes, *pk.meta_ptrs(), pk.B,
                                               out.data_ptr()), self.h)

    def wait(self) -> None:
        """Retires the oldest in-flight `submit_packed_host`."""
        _lib.check(self.lib.kdl_wait(self.h), self.h)

    def pack(self, images: Sequence[np.ndarray]) -> PackedImages:
        """Packs into this model's reusable pinned staging buffer (grown on demand).  The returned object is valid until the next
        `pack` call -- exactly what the synchronous request path needs."""
        need = sum((im.shape[0] * im.shape[1] * 3 + 255) // 256 * 256 for im in images) or 256
        st = getattr(self, "_staging", None)
        if st is None or st.numel() < need:
            self._staging = torch.empty(int(need * 1.25) + 256, dtype=torch.uint8, pin_memory=True)
        return PackedImages(images, staging=self._staging)

    def predict_uint8(self, images: Sequence[np.ndarray]) -> np.ndarray:
        outs = []
        for i in range(0, len(images), self.max_batch):
            outs.append(self.predict_packed_host(self.pack(images[i:i + self.max_batch])))
        return np.concatenate(outs, 0) if outs else np.empty((0, 10), np.float32)

    def profile_begin(self):
        """Bracket every kernel launch of this model with CUDA events (on the launching stream) until profile_end."""
        _lib.check(self.lib.kdl_profile_begin(self.h), self.h)

    def profile_end(self) -> List[dict]:
        rows = (_lib.KdlProfileRow * 16)()
        n = ctypes.c_int32()
        _lib.check(self.lib.kdl_profile_end(self.h, rows, 16, ctypes.byref(n)), self.h)
        return [{"kernel": rows[i].name.decode(), "launches": rows[i].launches, "ms": float(rows[i].ms),
                 "flops": rows[i].flops, "bytes": rows[i].bytes} for i in range(n.value)]

    # ------------------------------------------------------------------------------------------
    def tap_names(self) -> List[str]:
        return [self.lib.kdl_tap_name(self.h, i).decode() for i in range(self.lib.kdl_num_taps(self.h))]

    def layer_output(self, name: str):
        """(float32 numpy NHWC, relu_applied) of an intermediate tensor of the last forward (keep_taps=True)."""
        shape = (ctypes.c_int32 * 4)()
        relu = ctypes.c_int32()
        _lib.check(self.lib.kdl_layer_output(self.h, name.encode(), None, 0, ctypes.byref(shape), ctypes.byref(relu), None), self.h)
        n = int(np.prod(list(shape)))
        buf = torch.empty(n, dtype=torch.float32, device=self._dev())
        with torch.cuda.device(self.device):
            _lib.check(self.lib.kdl_layer_output(self.h, name.encode(), buf.data_ptr(), n, ctypes.byref(shape),
                                                 ctypes.byref(relu), self._stream()), self.h)
        torch.cuda.current_stream(self.device).synchronize()
        return buf.cpu().numpy().reshape(list(shape)), bool(relu.value)


def gemm_bf16(A: torch.Tensor, Wt: torch.Tensor, bias: torch.Tensor, residual: Optional[torch.Tensor] = None,
              relu: bool = False, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Stand-alone tcgen05 GEMM (the model's k_gemm_tc): act(A[M,K] @ Wt[N,K]^T + bias) (+ residual); bf16 CUDA tensors."""
    lib = _lib.load()
    M, K = A.shape
    N = Wt.shape[0]
    assert A.dtype == torch.bfloat16 and Wt.dtype == torch.bfloat16 and bias.dtype == torch.float32
    assert A.is_contiguous() and Wt.is_contiguous() and Wt.shape[1] == K
    if out is None:
        out = torch.empty((M, N), dtype=torch.bfloat16, device=A.device)
    st = ctypes.c_void_p(torch.cuda.current_stream(A.device).cuda_stream)
    with torch.cuda.device(A.device):
        _lib.check(lib.kdl_gemm_bf16(A.data_ptr(), Wt.data_ptr(), bias.data_ptr(),
                                     residual.data_ptr() if residual is not None else None, out.data_ptr(), M, N, K,
                                     int(relu), st))
    return out
