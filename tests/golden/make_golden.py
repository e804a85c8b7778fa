"""Generates the committed fixtures under tests/golden/.  Run HERE (the build container), where
/root/reference is mounted; the GPU box has no /root/reference, so tests only read the outputs.

  python tests/golden/make_golden.py

1. image2_fixture.npz  -- decoded pixels of the reference's only in-tree picture, media/image2.png
   (guide.md:610-612), plus what the reference's real preprocessing code (Pillow NEAREST resize + numpy
   fp32 scaling = the calls behind model_server.py:18,53) produces for it HERE: sha256 of the resized
   uint8 and of the (1,299,299,3) float32 tensor, spot values and the index tables.
2. resize_tables.npz    -- Pillow-generated index tables for a sweep of (n_in, n_out) pairs (known answers
   produced by the reference implementation itself, not by the restatement).
3. xception_golden.npz  -- fp64 oracle logits + tap statistics for seeded inputs/weights.  The network has
   no reference-side golden ("parity unpinned"), so these are regression pins of the restatement only.
"""
import hashlib
import json
import os
import sys

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def pillow_table(n_in, n_out):
    idx = np.arange(n_in, dtype=np.int32)
    img = np.stack([idx & 255, (idx >> 8) & 255, np.zeros_like(idx)], -1).astype(np.uint8)[None]
    r = np.asarray(Image.fromarray(img).resize((n_out, 1), Image.NEAREST))[0]
    return r[:, 0].astype(np.int32) | (r[:, 1].astype(np.int32) << 8)


def main():
    import PIL
    src = np.array(Image.open("/root/reference/media/image2.png").convert("RGB"))
    u8 = np.asarray(Image.fromarray(src).resize((299, 299), Image.NEAREST))
    x = np.array(u8, dtype="float32")[None]
    x /= 127.5
    x -= 1.0
    meta = {
        "source": "media/image2.png (400x534 RGB)", "pillow_version": PIL.__version__,
        "src_sha256": sha(src), "u8_sha256": sha(u8), "f32_sha256": sha(x),
        "u8_0_0": u8[0, 0].tolist(), "u8_149_149": u8[149, 149].tolist(), "u8_298_298": u8[298, 298].tolist(),
        "x_149_149": [float(v) for v in x[0, 149, 149]],
        "xtab": pillow_table(400, 299).tolist(), "ytab": pillow_table(534, 299).tolist(),
    }
    np.savez_compressed(os.path.join(HERE, "image2_fixture.npz"), src=src)
    with open(os.path.join(HERE, "image2_fixture.json"), "w") as f:
        json.dump(meta, f, indent=1)

    sizes = sorted(set(list(range(1, 64)) + [97, 149, 150, 151, 298, 299, 300, 301, 400, 534, 597, 598, 599, 640, 699,
                                             768, 1000, 1024, 1080, 1920, 2048, 4000, 8000]))
    tabs = {}
    for n_in in sizes:
        for n_out in (299, 150):
            tabs[f"{n_in}_{n_out}"] = pillow_table(n_in, n_out).astype(np.int32)
    np.savez_compressed(os.path.join(HERE, "resize_tables.npz"), **tabs)

    # network regression pins (fp64 oracle)
    import torch
    import kdl_b200  # noqa: F401
    from kdl_b200 import weights
    from oracle import xception_oracle as xo
    out = {}
    for h in (299, 150):
        rng = np.random.default_rng(0)
        u = rng.integers(0, 256, (2, h, h, 3), dtype=np.uint8)
        xin = u.astype(np.float32)
        xin /= 127.5
        xin -= 1.0
        for init in ("keras_default", "calibrated"):
            W = weights.make_weights(1, init)
            taps = {}
            y = xo.forward(W, xin, torch.float64, taps)
            out[f"logits_{h}_{init}"] = y
            for k, v in taps.items():
                out[f"tapstat_{h}_{init}_{k}"] = np.array([np.abs(v).mean(), v.mean(), v.reshape(-1)[:4].sum()])
    np.savez_compressed(os.path.join(HERE, "xception_golden.npz"), **out)
    print("ok", meta["u8_sha256"][:12], meta["f32_sha256"][:12], len(tabs), len(out))


if __name__ == "__main__":
    main()
