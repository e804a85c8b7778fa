"""Generates tests/golden/accuracy512.npz: fp64 oracle results for the 512-picture accuracy sample (SURVEY.md §8d config 1,
"accuracy over >= 512 images"), so that the GPU box only has to run the GPU side (the fp64 oracle needs ~1 s per picture).

  python tests/golden/make_accuracy_golden.py        # ~20 min on 8 cores

Per init (calibrated, calibrated_mild; weights seed 1): pictures = oracle.structured.structured_images(512, seed=3) with
mixed source sizes, preprocessed by Pillow (the reference's code path), pushed through oracle.xception_oracle in float64.
Stored: the natural-head logits, the re-centred head biases (oracle.structured.recentre_head on all 512 feature vectors) and
the logits under that head.  tests/test_gpu_accuracy.py compares the GPU against these.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

N, SEED = 512, 3


def main():
    import kdl_b200  # noqa: F401
    from kdl_b200 import weights
    from oracle import preprocess_oracle as po
    from oracle import structured as S
    from oracle import xception_oracle as xo
    imgs = S.structured_images(N, seed=SEED)
    out = {"n": np.array(N), "seed": np.array(SEED)}
    for init in ("calibrated", "calibrated_mild"):
        W = weights.make_weights(1, init)
        feats, logits = [], []
        for i in range(0, N, 16):
            x = np.concatenate([po.preprocess_pillow(im) for im in imgs[i:i + 16]])
            taps = {}
            logits.append(xo.forward(W, x, torch.float64, taps))
            feats.append(taps["avg_pool"])
            print(init, i, flush=True)
        feats = np.concatenate(feats)
        W2 = S.recentre_head(W, feats)
        out[f"{init}/logits_plain"] = np.concatenate(logits)
        out[f"{init}/dense_inner_bias"] = W2["dense_inner/bias"]
        out[f"{init}/dense_7_bias"] = W2["dense_7/bias"]
        out[f"{init}/logits_centred"] = S.head_only(W2, feats)
    np.savez_compressed(os.path.join(HERE, "accuracy512.npz"), **out)
    print("ok")


if __name__ == "__main__":
    main()
