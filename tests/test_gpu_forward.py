"""GPU parity of the Xception forward (through the C ABI) against the fp64 oracle.

Tolerances (BASELINE.json north_star): rel = max|a-b| / max|b| per image;
  fp32 path <= 1e-4;   bf16 path <= 2e-2 with identical top-1.  Layer-wise as well, because a logits-only check is blind
  to the middle flow under Keras-default init.

Inputs are structured pictures (tests/gpu_util.py) and the head biases are re-centred on the batch
(oracle.structured.recentre_head), so the logits are dominated by what distinguishes the pictures and the batch spans
several argmax classes: the tests assert that the fp64 oracle itself sees >= 4 different classes, which makes
"identical top-1" a statement about the image-dependent signal.

What the bf16 path is held to, per init (measured numbers: BASELINE.md §5):
  * keras_default (the reference's own random init) and calibrated_mild (every layer matters, residual branches are
    corrections to the skip path as in a trained network): the north_star contract literally -- every image <= 2e-2 on the
    init's natural head, identical top-1.
  * every init, re-centred head: identical top-1 on every picture whose fp64 top-2 margin exceeds the bf16 error bound, and
    the error of the image-dependent part <= 2e-2 of the batch's signal amplitude (gpu_util.batch_signal_rel).
  * calibrated (deliberately hostile: the residual stream doubles in variance per block, so every rounding error reaches
    the logits undamped; oracle/bf16_model.py reproduces its error on the CPU and shows that bf16 WEIGHT rounding is 85 % of
    it): mean <= 2e-2 and every image <= 3e-2 on the natural head -- the accepted deviation, recorded with its p95 / max."""
import numpy as np
import pytest
import torch

from gpu_util import BF16_TOL, F32_TOL, batch_signal_rel, rel, synth_inputs, top2_margin

pytestmark = pytest.mark.gpu


def _oracle(W, x, taps=None):
    from oracle import xception_oracle as xo
    return xo.forward(W, x, torch.float64, taps)


def _mk(W, **kw):
    import kdl_b200  # noqa: F401
    from kdl_b200 import engine
    return engine.Model(W, **kw)


@pytest.mark.parametrize("h", [299, 150])
@pytest.mark.parametrize("init", ["calibrated", "keras_default"])
def test_fp32_path_logits_and_layers(h, init):
    from kdl_b200 import weights
    W = weights.make_weights(1, init)
    _, x = synth_inputs(h, 3)
    taps = {}
    want = _oracle(W, x, taps)
    m = _mk(W, max_batch=4, input_size=h, precision="f32", keep_taps=True)
    try:
        got = m.serving_default(input_8=x)["dense_7"]
        assert got.shape == (3, 10) and got.dtype == np.float32
        for i in range(3):
            assert rel(got[i], want[i]) <= F32_TOL, (i, got[i], want[i])
        names = m.tap_names()
        assert len(names) == 45
        for name in names:
            a, relu = m.layer_output(name)
            b = taps[name]
            if relu:
                b = np.maximum(b, 0)
            if b.ndim == 2:
                b = b.reshape(a.shape)
            assert a.shape == b.shape, name
            assert rel(a, b) <= F32_TOL, (name, rel(a, b))
    finally:
        m.close()


@pytest.mark.parametrize("gemm", ["simple", "auto"])
@pytest.mark.parametrize("h,init", [(299, "calibrated"), (150, "calibrated"), (299, "keras_default"), (299, "calibrated_mild"),
                                    (150, "calibrated_mild")])
def test_bf16_path_logits_top1_and_layers(h, init, gemm):
    from kdl_b200 import weights
    from oracle import structured as S
    W = weights.make_weights(1, init)
    n = 16
    _, x = synth_inputs(h, n)
    taps = {}
    want = _oracle(W, x, taps)
    # the same network with the head re-centred on this batch: the discriminating variant
    Wc = S.recentre_head(W, taps["avg_pool"])
    want_c = S.head_only(Wc, taps["avg_pool"])
    if init != "keras_default":                      # (Keras-default logits are ~1e-5: its argmax is the same for every picture)
        assert len(set(want_c.argmax(1).tolist())) >= 4, want_c.argmax(1)

    m = _mk(W, max_batch=n, input_size=h, precision="bf16", gemm_impl=gemm, keep_taps=True)
    try:
        got = m.serving_default(input_8=x)["dense_7"]
        errs = np.array([rel(got[i], want[i]) for i in range(n)])
        if init == "calibrated":                     # hostile init: the accepted deviation
            assert float(errs.mean()) <= BF16_TOL and float(errs.max()) <= 1.5 * BF16_TOL, errs
        else:                                        # the contract, literally
            assert float(errs.max()) <= BF16_TOL, errs
        # identical top-1 wherever the fp64 top-2 margin exceeds the error the contract allows (a tie within the tolerance can
        # legitimately fall either way; on this batch every picture is decided except under the ~1e-5 Keras-default logits)
        dec_nat = top2_margin(want) > 2 * BF16_TOL * np.abs(want).max(1)
        assert (got.argmax(1)[dec_nat] == want.argmax(1)[dec_nat]).all(), (got.argmax(1), want.argmax(1))
        worst = 0.0
        for name in m.tap_names():
            a, relu = m.layer_output(name)
            b = taps[name]
            if relu:
                b = np.maximum(b, 0)
            b = b.reshape(a.shape)
            e = rel(a, b)
            worst = max(worst, e)
            # per-layer drift stays within the end-to-end bf16 budget (keras_default branches are ~1e-6 of the
            # skip path, so only their block outputs are meaningful there)
            if init != "keras_default" or name.endswith("_out") or name.startswith("block1"):
                assert e <= 1.5 * BF16_TOL, (name, e)
    finally:
        m.close()

    mc = _mk(Wc, max_batch=n, input_size=h, precision="bf16", gemm_impl=gemm)
    try:
        got_c = mc.serving_default(input_8=x)["dense_7"]
    finally:
        mc.close()
    sig = batch_signal_rel(got_c, want_c)
    assert float(sig.max()) <= (1.5 * BF16_TOL if init == "calibrated" else BF16_TOL), sig
    # identical top-1 wherever the fp64 margin is larger than the error the path is allowed to have
    bound = BF16_TOL * np.abs(want_c - want_c.mean(0, keepdims=True)).max()
    decided = top2_margin(want_c) > 2 * bound
    if init != "keras_default":
        assert decided.sum() >= n // 2, top2_margin(want_c)
    assert (got_c.argmax(1)[decided] == want_c.argmax(1)[decided]).all(), (got_c.argmax(1), want_c.argmax(1))
    print(f"bf16 {gemm} h={h} {init}: natural head rel mean {errs.mean():.2e} max {errs.max():.2e}; re-centred head signal err "
          f"max {sig.max():.2e}, top-1 {got_c.argmax(1).tolist()} ({int(decided.sum())}/{n} decided); worst layer {worst:.2e}")


def test_fused_uint8_path_matches_two_step():
    """predict_uint8 (resize+scale fused into the conv1 load) == preprocess then serving_default: bit for bit when both run the
    CUDA-core conv1 (fp32 path; bf16 path with KdlTuning.conv1_tc = 0).  With the tensor-core conv1 (the bf16 default for uint8
    input: raw pixel values are exact in bf16, x/127.5-1 is folded into bf16 weights) only conv1's weight rounding differs."""
    from kdl_b200 import weights
    from oracle import structured as S
    W = weights.make_weights(1, "calibrated")
    imgs = S.structured_images(5)
    for prec, tuning in (("f32", None), ("bf16", {"conv1_tc": 0}), ("bf16", None)):
        m = _mk(W, max_batch=8, precision=prec, tuning=tuning)
        try:
            x = m.preprocess_uint8(imgs)
            two = m.forward_device(x).cpu().numpy()
            one = m.predict_uint8(imgs)
            assert one.shape == (5, 10)
            if prec == "f32" or tuning:
                np.testing.assert_array_equal(one, two)   # same arithmetic, same kernels after conv1's load
            else:
                errs = [rel(one[i], two[i]) for i in range(5)]
                assert 0 < max(errs) <= 1e-2, errs
        finally:
            m.close()


def test_tensor_core_conv1_against_the_oracle():
    """block1_conv1 as a tcgen05 implicit GEMM on raw uint8 pixels (conv1_tc.cuh), checked where it writes: the block1_conv1_act
    tap of the uint8 path against the fp64 oracle on the Pillow-preprocessed pictures, mixed source sizes, two batch sizes (a
    partial last tile and >1 tile per CTA), and against the CUDA-core conv1 (fp32 weights) on the same input."""
    from kdl_b200 import weights
    from oracle import preprocess_oracle as po
    from oracle import structured as S
    W = weights.make_weights(1, "calibrated")
    for n, size in ((3, 299), (7, 150)):
        imgs = S.structured_images(n)
        x = np.concatenate([po.preprocess_pillow(im, size, size) for im in imgs])
        taps = {}
        _oracle(W, x[:2], taps)
        want = np.maximum(taps["block1_conv1_act"], 0)
        got = {}
        for tc in (1, 0):
            m = _mk(W, max_batch=8, input_size=size, precision="bf16", keep_taps=True, tuning={"conv1_tc": tc})
            try:
                m.predict_uint8(imgs)
                got[tc], relu = m.layer_output("block1_conv1_act")
                assert relu and got[tc].shape[0] == n
            finally:
                m.close()
        assert rel(got[1][:2], want) <= 8e-3, rel(got[1][:2], want)        # bf16 output rounding (4e-3) + bf16 weights
        assert rel(got[0][:2], want) <= 5e-3
        assert rel(got[1], got[0]) <= 8e-3 and not np.array_equal(got[1], got[0])


def test_batch_split_and_head_variant():
    from kdl_b200 import weights
    from oracle import xception_oracle as xo
    W = weights.make_weights(3, "calibrated", head_inner=0)
    _, x = synth_inputs(150, 5, seed=5)
    want = xo.forward(W, x, torch.float64)
    m = _mk(W, max_batch=2, input_size=150, precision="f32")     # 5 images through max_batch 2 -> 3 calls
    try:
        got = m.serving_default(input_8=x)["dense_7"]
        for i in range(5):
            assert rel(got[i], want[i]) <= F32_TOL
        with pytest.raises(KeyError):
            m.serving_default(images=x)
        with pytest.raises(ValueError):
            m.serving_default(input_8=x[:, :100])
    finally:
        m.close()
