"""Pins the preprocessing oracle (oracle/preprocess_oracle.{c,py}) against the reference's real
implementation (Pillow NEAREST + numpy fp32 scaling: model_server.py:18,53) and the committed
known answers.  CPU only."""
import hashlib
import json
import os

import numpy as np
import pytest

from oracle import preprocess_oracle as po


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.fixture(scope="module")
def fixture(golden_dir):
    src = np.load(os.path.join(golden_dir, "image2_fixture.npz"))["src"]
    meta = json.load(open(os.path.join(golden_dir, "image2_fixture.json")))
    return src, meta


def test_fixture_source_hash(fixture):
    src, meta = fixture
    assert src.shape == (534, 400, 3)
    # SURVEY.md §8c (i)
    assert meta["src_sha256"].startswith("db048dd86f21aabb")
    assert sha(src) == meta["src_sha256"]


@pytest.mark.parametrize("fn", ["resize_nearest_np", "resize_c", "resize_pillow"])
def test_image2_resize_known_answer(fixture, fn):
    src, meta = fixture
    u8 = getattr(po, fn)(src)
    assert sha(u8) == "91075d2a131a0f89b86c3156f7f3715f445454602b1c8e65e43a96002fae3848" == meta["u8_sha256"]
    assert u8[0, 0].tolist() == [113, 108, 102]
    assert u8[149, 149].tolist() == [237, 232, 226]
    assert u8[298, 298].tolist() == [53, 24, 10]


@pytest.mark.parametrize("fn", ["preprocess_np", "preprocess_c", "preprocess_pillow"])
def test_image2_preprocess_known_answer(fixture, fn):
    src, meta = fixture
    x = getattr(po, fn)(src)
    assert x.shape == (1, 299, 299, 3) and x.dtype == np.float32
    assert sha(x) == "3c14452b6dcf1eeb5b8f0475127f980b15b20440ae7eb64b9ede18cc8098aa67" == meta["f32_sha256"]
    np.testing.assert_allclose(x[0, 149, 149], [0.8588235378, 0.8196078539, 0.7725490332], rtol=0, atol=1e-9)


def test_image2_index_tables(fixture):
    _, meta = fixture
    xt, yt = po.index_table(400, 299), po.index_table(534, 299)
    assert xt[:8].tolist() == [0, 2, 3, 4, 6, 7, 8, 10] and xt[-4:].tolist() == [395, 396, 397, 399]
    assert yt[:8].tolist() == [0, 2, 4, 6, 8, 9, 11, 13] and yt[-4:].tolist() == [527, 529, 531, 533]
    assert xt.tolist() == meta["xtab"] and yt.tolist() == meta["ytab"]
    assert po.index_table_c(400, 299).tolist() == meta["xtab"]


def test_tables_against_pillow_generated_goldens(golden_dir):
    tabs = np.load(os.path.join(golden_dir, "resize_tables.npz"))
    assert len(tabs.files) >= 150
    for key in tabs.files:
        n_in, n_out = map(int, key.split("_"))
        want = tabs[key]
        assert np.array_equal(po.index_table(n_in, n_out), want), key
        assert np.array_equal(po.index_table_c(n_in, n_out), want), key


def test_tables_against_live_pillow_sweep():
    from PIL import Image
    for n_in in list(range(1, 700, 7)) + [299, 598, 2048, 4097]:
        for n_out in (299, 150):
            idx = np.arange(n_in, dtype=np.int32)
            img = np.stack([idx & 255, (idx >> 8) & 255, 0 * idx], -1).astype(np.uint8)[None]
            r = np.asarray(Image.fromarray(img).resize((n_out, 1), Image.NEAREST))[0]
            got = r[:, 0].astype(np.int32) | (r[:, 1].astype(np.int32) << 8)
            assert np.array_equal(got, po.index_table_c(n_in, n_out)), (n_in, n_out)


def test_closed_form_is_not_the_rule():
    """floor((x+0.5)*in/out) is NOT what Pillow does; the oracle must not use it."""
    diff = 0
    for n_in in range(1, 700):
        cf = np.floor((np.arange(299) + 0.5) * n_in / 299).astype(np.int32)
        diff += int(not np.array_equal(cf, po.index_table(n_in, 299)))
    assert diff > 0


@pytest.mark.parametrize("h,w", [(1, 1), (1, 7), (7, 1), (2, 3), (299, 299), (150, 150), (300, 298), (64, 2048), (1080, 1920)])
def test_random_images_all_paths_agree(h, w):
    rng = np.random.default_rng(h * 10007 + w)
    img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    for oh in (299, 150):
        want = po.resize_pillow(img, oh, oh)
        assert np.array_equal(po.resize_nearest_np(img, oh, oh), want)
        assert np.array_equal(po.resize_c(img, oh, oh), want)
        x = po.preprocess_pillow(img, oh, oh)
        assert x.tobytes() == po.preprocess_c(img, oh, oh).tobytes() == po.preprocess_np(img, oh, oh).tobytes()


def test_identity_size_is_a_copy():
    rng = np.random.default_rng(3)
    img = rng.integers(0, 256, (299, 299, 3), dtype=np.uint8)
    assert np.array_equal(po.resize_c(img), img)


def test_scale_two_roundings_all_256_values():
    """x/127.5 then -1 in fp32 (two roundings); x*(1/127.5f)-1 differs on many inputs."""
    v = np.arange(256, dtype=np.uint8).reshape(1, 256, 1).repeat(3, 2)
    want = po.scale_np(v)[0, 0, :, 0]
    got = po.preprocess_c(np.ascontiguousarray(v), 1, 256)[0, 0, :, 0]
    assert want.tobytes() == got.tobytes()
    mul = np.arange(256, dtype=np.float32) * np.float32(1 / 127.5) - np.float32(1.0)
    assert (mul != want).sum() > 50
    assert want.min() == -1.0 and want.max() == 1.0


def test_rgba_and_gray_are_converted_like_the_gateway():
    from PIL import Image
    rng = np.random.default_rng(5)
    rgba = rng.integers(0, 256, (40, 50, 4), dtype=np.uint8)
    rgb = np.asarray(Image.fromarray(rgba, "RGBA").convert("RGB"))
    assert np.array_equal(rgb, rgba[..., :3])  # alpha dropped
