"""accuracy of the bf16 path over >= 512 pictures at batch 64, against fp64 oracle results that
were computed once in the build container (tests/golden/accuracy512.npz, made by tests/golden/make_accuracy_golden.py --
the fp64 oracle needs ~1 s per picture, the GPU box should not spend minutes on it).  The pictures are regenerated here from
the same seeds (oracle.structured.structured_images: mixed source sizes, so the resize kernel is part of what is checked).

Asserted (and written to gpurun_out/accuracy512.json for BASELINE.md):
  * calibrated_mild, natural head: the north_star contract -- EVERY one of the 512 pictures within 2e-2; top-1 identical on every
    picture whose fp64 top-2 margin is larger than the error the contract allows (2 x 2e-2 x max|logit|), >= 99.5 % overall (a
    picture whose two best classes are tied to within the tolerance can legitimately land on either side);
  * calibrated (hostile init), natural head: mean <= 2e-2, p95 <= 2.5e-2, max <= 3.5e-2, top-1 as above with >= 95 % overall --
    the accepted deviation (oracle/bf16_model.py predicts mean 1.1e-2 / p95 1.7e-2 / max 2.6e-2 from bf16 rounding alone);
  * both inits, head re-centred on the 512 feature vectors (>= 6 argmax classes): top-1 agreement on every picture whose fp64
    margin exceeds the error bound, >= 99 % overall, and the image-dependent error <= 2e-2 (mild) / 3e-2 (hostile) of the
    signal amplitude."""
import json
import os

import numpy as np
import pytest

from gpu_util import BF16_TOL, batch_signal_rel, rel, top2_margin

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("init", ["calibrated_mild", "calibrated"])
def test_bf16_accuracy_over_512_structured_pictures(golden_dir, init):
    import kdl_b200  # noqa: F401
    from kdl_b200 import engine, weights
    from oracle import structured as S
    g = np.load(os.path.join(golden_dir, "accuracy512.npz"))
    n = int(g["n"])
    assert n >= 512
    imgs = S.structured_images(n, seed=int(g["seed"]))
    W = weights.make_weights(1, init)
    Wc = dict(W)
    Wc["dense_inner/bias"] = g[f"{init}/dense_inner_bias"]
    Wc["dense_7/bias"] = g[f"{init}/dense_7_bias"]
    want, want_c = g[f"{init}/logits_plain"], g[f"{init}/logits_centred"]
    assert len(set(want_c.argmax(1).tolist())) >= 6

    def run(weights_):
        m = engine.Model(weights_, max_batch=64, precision="bf16")
        try:
            return m.predict_uint8(imgs)                      # 8 calls of 64
        finally:
            m.close()

    got, got_c = run(W), run(Wc)
    errs = np.array([rel(got[i], want[i]) for i in range(n)])
    sig = batch_signal_rel(got_c, want_c)
    bound = BF16_TOL * np.abs(want_c - want_c.mean(0, keepdims=True)).max()
    decided = top2_margin(want_c) > 2 * bound
    agree_c = got_c.argmax(1) == want_c.argmax(1)
    agree = got.argmax(1) == want.argmax(1)
    decided_nat = top2_margin(want) > 2 * BF16_TOL * np.abs(want).max(1)
    rep = {"init": init, "images": n, "batch": 64,
           "natural_head": {"rel_mean": float(errs.mean()), "rel_p50": float(np.median(errs)), "rel_p95": float(np.percentile(errs, 95)),
                            "rel_max": float(errs.max()), "frac_within_2e-2": float((errs <= BF16_TOL).mean()),
                            "top1_agreement": float(agree.mean()), "decided": int(decided_nat.sum()),
                            "top1_agreement_decided": float(agree[decided_nat].mean())},
           "recentred_head": {"classes_in_oracle_top1": len(set(want_c.argmax(1).tolist())), "signal_err_p95": float(np.percentile(sig, 95)),
                              "signal_err_max": float(sig.max()), "top1_agreement": float(agree_c.mean()),
                              "decided": int(decided.sum()), "top1_agreement_decided": float(agree_c[decided].mean())}}
    os.makedirs("gpurun_out", exist_ok=True)
    with open(os.path.join("gpurun_out", f"accuracy512_{init}.json"), "w") as f:
        json.dump(rep, f, indent=1)
    print(json.dumps(rep))
    hostile = init == "calibrated"
    if not hostile:
        assert errs.max() <= BF16_TOL, rep
    else:
        assert errs.mean() <= BF16_TOL and np.percentile(errs, 95) <= 2.5e-2 and errs.max() <= 3.5e-2, rep
    assert agree[decided_nat].all() and decided_nat.sum() >= n // 2, rep
    assert agree.mean() >= (0.95 if hostile else 0.995), rep
    assert sig.max() <= (1.5 * BF16_TOL if hostile else BF16_TOL), rep
    assert agree_c[decided].all(), rep
    assert agree_c.mean() >= 0.99, rep
