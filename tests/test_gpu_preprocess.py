"""GPU parity of K1 (fused nearest-resize + scale) against the preprocessing oracle: BIT-EXACT."""
import hashlib
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def model():
    import kdl_b200  # noqa: F401
    from kdl_b200 import engine, weights
    m = engine.Model(weights.make_weights(1, "calibrated"), max_batch=16, precision="f32")
    yield m
    m.close()


@pytest.fixture(scope="module")
def model150():
    import kdl_b200  # noqa: F401
    from kdl_b200 import engine, weights
    m = engine.Model(weights.make_weights(1, "calibrated"), max_batch=8, input_size=150, precision="f32")
    yield m
    m.close()


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_image2_known_answer(model, golden_dir):
    src = np.load(os.path.join(golden_dir, "image2_fixture.npz"))["src"]
    meta = json.load(open(os.path.join(golden_dir, "image2_fixture.json")))
    x = model.preprocess_uint8([src]).cpu().numpy()
    assert x.shape == (1, 299, 299, 3) and x.dtype == np.float32
    assert sha(x) == meta["f32_sha256"] == "3c14452b6dcf1eeb5b8f0475127f980b15b20440ae7eb64b9ede18cc8098aa67"


@pytest.mark.parametrize("sizes", [
    [(1, 1)], [(1, 7), (7, 1), (2, 3)], [(299, 299), (298, 300), (300, 298)], [(534, 400), (64, 2048), (2048, 64)],
    [(1080, 1920), (150, 150), (149, 151), (597, 598), (33, 1000), (1000, 33), (299, 598), (640, 480)],
])
def test_mixed_sizes_bit_exact(model, sizes):
    from oracle import preprocess_oracle as po
    rng = np.random.default_rng(len(sizes) * 7919 + sizes[0][0])
    imgs = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for h, w in sizes]
    got = model.preprocess_uint8(imgs).cpu().numpy()
    for i, im in enumerate(imgs):
        want = po.preprocess_c(im)[0]
        assert got[i].tobytes() == want.tobytes(), sizes[i]
        assert want.tobytes() == po.preprocess_pillow(im)[0].tobytes()


def test_random_size_sweep_bit_exact(model):
    from oracle import preprocess_oracle as po
    rng = np.random.default_rng(2)
    for _ in range(12):
        imgs = [rng.integers(0, 256, (int(rng.integers(1, 700)), int(rng.integers(1, 700)), 3), dtype=np.uint8)
                for _ in range(16)]
        got = model.preprocess_uint8(imgs).cpu().numpy()
        for i, im in enumerate(imgs):
            assert got[i].tobytes() == po.preprocess_c(im)[0].tobytes(), im.shape


def test_150_target(model150):
    from oracle import preprocess_oracle as po
    rng = np.random.default_rng(4)
    imgs = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for h, w in [(534, 400), (150, 150), (75, 301), (1, 1)]]
    got = model150.preprocess_uint8(imgs).cpu().numpy()
    for i, im in enumerate(imgs):
        assert got[i].tobytes() == po.preprocess_pillow(im, 150, 150)[0].tobytes()


def test_all_256_values_scale_exactly(model):
    from oracle import preprocess_oracle as po
    img = np.arange(256, dtype=np.uint8).reshape(1, 256, 1).repeat(3, 2).repeat(4, 0)
    got = model.preprocess_uint8([np.ascontiguousarray(img)]).cpu().numpy()
    assert got.tobytes() == po.preprocess_np(np.ascontiguousarray(img)).tobytes()


def test_row_pitch_and_bad_geometry(model):
    import ctypes
    import torch
    from kdl_b200 import _lib
    from oracle import preprocess_oracle as po
    rng = np.random.default_rng(9)
    full = rng.integers(0, 256, (50, 80, 3), dtype=np.uint8)
    view = full[5:45, 10:70]                       # pitch 240 B, width 60
    d = torch.from_numpy(full.reshape(-1)).cuda()
    off = np.array([5 * 240 + 10 * 3], np.int64); h = np.array([40], np.int32); w = np.array([60], np.int32)
    pitch = np.array([240], np.int64)
    out = torch.empty((1, 299, 299, 3), dtype=torch.float32, device="cuda")
    rc = model.lib.kdl_preprocess_u8(model.h, d.data_ptr(), off.ctypes.data, h.ctypes.data, w.ctypes.data,
                                     pitch.ctypes.data, 1, out.data_ptr(), None)
    assert rc == 0
    torch.cuda.synchronize()
    assert out.cpu().numpy().tobytes() == po.preprocess_c(np.ascontiguousarray(view)).tobytes()
    pitch[0] = 10                                    # pitch < w*3 -> rejected, with a message
    rc = model.lib.kdl_preprocess_u8(model.h, d.data_ptr(), off.ctypes.data, h.ctypes.data, w.ctypes.data,
                                     pitch.ctypes.data, 1, out.data_ptr(), None)
    assert rc == _lib.KDL_EINVAL and b"bad geometry" in model.lib.kdl_last_error(model.h)
    rc = model.lib.kdl_preprocess_u8(model.h, d.data_ptr(), off.ctypes.data, h.ctypes.data, w.ctypes.data,
                                     pitch.ctypes.data, 17, out.data_ptr(), None)
    assert rc == _lib.KDL_EINVAL                     # batch > max_batch
