"""Drop-in mirror of the reference gateway's model path (model_server.py), minus HTTP/gRPC transport.

Same names, argument meaning and error behaviour as model_server.py for the hot path:
  labels (21-32), create_preprocessor(name, target_size) (18), make_request / process_response (38-49),
  apply_model (52-56).  `apply_model(url)` still returns {label: float} over the 10 clothing classes in the
  reference's order; what changes is that `stub.Predict` is an in-process call into the CUDA engine.

Image download + decode (urllib + PIL decoders inside keras-image-helper's from_url) stays on the CPU and is
out of scope; inputs to the accelerated path are decoded RGB uint8 arrays.
"""
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

    def __init__(self, model: Model, target_size=(299, 299)):
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
        return self.model.preprocess_uint8([rgb_u8]).cpu().numpy()

    def from_path(self, path: str) -> np.ndarray:
        with open(path, 'rb') as f:
            return self.from_array(self.decode(f.read()))

    def from_url(self, url: str) -> np.ndarray:
        with request.urlopen(url) as resp:
            buffer = resp.read()
        return self.from_array(self.decode(buffer))


def create_preprocessor(name: str, target_size, model: Model) -> Preprocessor:
    """model_server.py:18 `create_preprocessor('xception', target_size=(299, 299))`."""
    if name != 'xception':
        raise ValueError(f"only the 'xception' preprocessing recipe is served here, got {name!r}")
    return Preprocessor(model, target_size)


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

    def __init__(self, model: Model):
        self.model = model
        self.preprocessor = create_preprocessor('xception', (model.input_size, model.input_size), model)

    def Predict(self, pb_request: dict, timeout: Optional[float] = None) -> dict:
        spec = pb_request["model_spec"]
        if spec["name"] != MODEL_NAME:
            raise LookupError(f"Servable not found for request: Latest({spec['name']})")   # TF-Serving NOT_FOUND
        if spec["signature_name"] != SIGNATURE_NAME:
            raise LookupError(f"Serving signature key \"{spec['signature_name']}\" not found.")
        return {"outputs": self.model.serving_default(**pb_request["inputs"])}

    def apply_model_array(self, rgb_u8: np.ndarray) -> Dict[str, float]:
        """apply_model on an already decoded image: fused uint8 path (resize+scale+Xception in one call)."""
        logits = self.model.predict_uint8([rgb_u8])
        return {c: float(p) for c, p in zip(labels, logits[0])}

    def apply_model(self, url: str) -> Dict[str, float]:
        """model_server.py:52-56."""
        X = self.preprocessor.from_url(url)
        pb_request = make_request(X)
        pb_result = self.Predict(pb_request, timeout=20.0)
        return process_response(pb_result)

    def predict_batch(self, images: List[np.ndarray]) -> List[Dict[str, float]]:
        logits = self.model.predict_uint8(images)
        return [{c: float(p) for c, p in zip(labels, row)} for row in logits]
