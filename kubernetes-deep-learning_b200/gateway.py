"""Drop-in mirror of the reference gateway's model path (model_server.py), minus HTTP/gRPC transport.

Same names, argument meaning and error behaviour as model_server.py for the hot path:
  labels (21-32), create_preprocessor(name, target_size) (18), make_request / process_response (38-49),
  apply_model (52-56).  `apply_model(url)` still returns {label: float} over the 10 clothing classes in the
  reference's order; what changes is that `stub.Predict` is an in-process call into the CUDA engine.

Image download + decode (urllib + PIL decoders inside keras-image-helper's from_url) stays on the CPU and is
out of scope; inputs to the accelerated path are decoded RGB uint8 arrays.
"""
import contextlib
import threading
from io import BytesIO
from typing import Dict, List, Optional
from urllib import request

import numpy as np

from .engine import INPUT_KEY, MODEL_NAME, OUTPUT_KEY, SIGNATURE_NAME, Model

# model_server.py:21-32 -- order is part of the contract
labels = [
    'dress',
    'hat',
    'longsleeve',
    'outwear',
    'pants',
    'shirt',
    'shoes',
    'shorts',
    'skirt',
    't-shirt'
]


class Preprocessor:
    """keras-image-helper 0.0.1 preprocessor surface (`from_url`, `from_path`, `resize_image`,
    `convert_to_tensor`), with resize + scaling executed by the fused CUDA kernel."""

    def __init__(self, model: Model, target_size=(299, 299), lock=None):
        self._lock = lock if lock is not None else contextlib.nullcontext()   # shared with whoever else drives `model`
        if tuple(target_size) != (model.input_size, model.input_size):
            raise ValueError(f"target_size {target_size} does not match the model's input size {model.input_size}")
        self.model = model
        self.target_size = tuple(target_size)

    @staticmethod
    def decode(buffer: bytes) -> np.ndarray:
        from PIL import Image
        img = Image.open(BytesIO(buffer))
        if img.mode != 'RGB':
            img = img.convert('RGB')
        return np.asarray(img)

    def from_array(self, rgb_u8: np.ndarray) -> np.ndarray:
        """decoded RGB uint8 HWC -> float32 (1,S,S,3) in [-1,1] (host array, like the reference returns)."""
        with self._lock:
            return self.model.preprocess_uint8([rgb_u8]).cpu().numpy()

    def from_path(self, path: str) -> np.ndarray:
        with open(path, 'rb') as f:
            return self.from_array(self.decode(f.read()))

    def fetch(self, url: str) -> np.ndarray:
        """download + decode -> RGB uint8 HWC (the CPU part of from_url)"""
        with request.urlopen(url) as resp:
            buffer = resp.read()
        return self.decode(buffer)

    def from_url(self, url: str) -> np.ndarray:
        return self.from_array(self.fetch(url))


def create_preprocessor(name: str, target_size, model: Model, lock=None) -> Preprocessor:
    """model_server.py:18 `create_preprocessor('xception', target_size=(299, 299))`."""
    if name != 'xception':
        raise ValueError(f"only the 'xception' preprocessing recipe is served here, got {name!r}")
    return Preprocessor(model, target_size, lock)


def make_request(X: np.ndarray) -> dict:
    """model_server.py:38-43: the PredictRequest fields as a plain dict (no protobuf on the hot path)."""
    return {"model_spec": {"name": MODEL_NAME, "signature_name": SIGNATURE_NAME}, "inputs": {INPUT_KEY: X}}


def process_response(result: dict) -> Dict[str, float]:
    """model_server.py:46-49: zip(labels, outputs['dense_7'].float_val) -- a flat float list, so with a batch
    only image 0's ten logits are returned, exactly as the reference does."""
    pred = np.asarray(result["outputs"][OUTPUT_KEY]).reshape(-1).tolist()
    return {c: p for c, p in zip(labels, pred)}


class Gateway:
    """Module-level state of model_server.py (stub + preprocessor) as an object."""

    def __init__(self, model: Model, batching=None):
        """batching: a batching.BatchingParameters coalesces concurrent apply_model* calls (from the HTTP server's request
        threads) into one forward call; None = one request at a time, the reference's single sync worker (gateway.dockerfile:17)."""
        self.model = model
        self._lock = threading.Lock()     # a KdlModel is not re-entrant
        self.preprocessor = create_preprocessor('xception', (model.input_size, model.input_size), model, self._lock)
        self._u8_batcher = self._f32_batcher = None
        if batching is not None:
            from .batching import DynamicBatcher
            self._u8_batcher = DynamicBatcher(self._run_u8, batching)
            self._f32_batcher = DynamicBatcher(self._run_f32, batching)

    def _run_u8(self, images):
        with self._lock:
            return list(self.model.predict_uint8(list(images)))

    def _run_f32(self, xs):
        x = xs[0] if len(xs) == 1 else np.concatenate(xs, axis=0)
        with self._lock:
            y = self.model.serving_default(**{INPUT_KEY: x})[OUTPUT_KEY]
        out, o = [], 0
        for a in xs:
            out.append(y[o:o + a.shape[0]])
            o += a.shape[0]
        return out

    def close(self):
        for b in (self._u8_batcher, self._f32_batcher):
            if b is not None:
                b.close()

    def Predict(self, pb_request: dict, timeout: Optional[float] = None) -> dict:
        spec = pb_request["model_spec"]
        if spec["name"] != MODEL_NAME:
            raise LookupError(f"Servable not found for request: Latest({spec['name']})")   # TF-Serving NOT_FOUND
        if spec["signature_name"] != SIGNATURE_NAME:
            raise LookupError(f"Serving signature key \"{spec['signature_name']}\" not found.")
        inputs = pb_request["inputs"]
        if set(inputs) != {INPUT_KEY}:
            raise ValueError(f"input tensor alias not found in signature: expected exactly {{{INPUT_KEY}}}, got {sorted(inputs)}")
        x = inputs[INPUT_KEY]
        if self._f32_batcher is not None and x.shape[0] > 0:
            y = self._f32_batcher.submit(x, size=int(x.shape[0]), timeout=timeout)
        else:
            y = self._run_f32([x])[0]
        return {"outputs": {OUTPUT_KEY: y}}

    def apply_model_array(self, rgb_u8: np.ndarray) -> Dict[str, float]:
        """apply_model on an already decoded image: fused uint8 path (resize+scale+Xception in one call)."""
        if self._u8_batcher is not None:
            row = self._u8_batcher.submit(rgb_u8)
        else:
            row = self._run_u8([rgb_u8])[0]
        return {c: float(p) for c, p in zip(labels, row)}

    def apply_model(self, url: str) -> Dict[str, float]:
        """model_server.py:52-56.  With batching on, the decoded image goes straight to the fused uint8 path so that concurrent
        requests share one resize+forward call (same logits: tests/test_gpu_gateway.py)."""
        if self._u8_batcher is not None:
            return self.apply_model_array(self.preprocessor.fetch(url))
        X = self.preprocessor.from_url(url)
        pb_request = make_request(X)
        pb_result = self.Predict(pb_request, timeout=20.0)
        return process_response(pb_result)

    def predict_batch(self, images: List[np.ndarray]) -> List[Dict[str, float]]:
        logits = self._run_u8(images)
        return [{c: float(p) for c, p in zip(labels, row)} for row in logits]
