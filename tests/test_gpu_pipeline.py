"""Pipelined host path (kdl_submit_u8_host / kdl_wait): same logits as the synchronous call, in order, with two submits in flight."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_submit_wait_matches_synchronous_call():
    import kdl_b200  # noqa: F401
    from kdl_b200 import engine, weights
    W = weights.make_weights(1, "calibrated")
    m = engine.Model(W, max_batch=6, precision="bf16")
    try:
        rng = np.random.default_rng(3)
        batches = [[rng.integers(0, 256, (int(rng.integers(32, 700)), int(rng.integers(32, 700)), 3), dtype=np.uint8)
                    for _ in range(int(rng.integers(1, 7)))] for _ in range(7)]
        packs = [engine.PackedImages(b) for b in batches]
        want = [m.predict_packed_host(p).copy() for p in packs]
        outs = [torch.zeros((6, 10), dtype=torch.float32).pin_memory() for _ in packs]
        m.submit_packed_host(packs[0], outs[0])
        for i in range(1, len(packs)):
            m.submit_packed_host(packs[i], outs[i])      # second in flight
            m.wait()                                     # retires i-1
            np.testing.assert_array_equal(outs[i - 1][:packs[i - 1].B].numpy(), want[i - 1])
        m.wait()
        np.testing.assert_array_equal(outs[-1][:packs[-1].B].numpy(), want[-1])
        m.wait()                                         # nothing in flight: a no-op
        for i in range(3):                               # a third submit retires the oldest by itself
            m.submit_packed_host(packs[i], outs[i])
        m.wait(); m.wait()
        for i in range(3):
            np.testing.assert_array_equal(outs[i][:packs[i].B].numpy(), want[i])
    finally:
        m.close()
