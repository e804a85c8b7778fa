"""Real-weights importer (keras_io.py): Keras HDF5 / npz layouts of the clothing model -> weight bundle.  CPU only.
The trained .h5 (convert.py:4) is unreachable offline, so the files are written here in the layout Keras uses."""
import numpy as np
import pytest

import kdl_b200  # noqa: F401
from kdl_b200 import arch, hdf5_min, keras_io, weights


@pytest.fixture(scope="module")
def W():
    return weights.make_weights(3, "calibrated")


def _same(a, b):
    assert list(a) == list(b)
    for k in a:
        assert a[k].dtype == np.float32
        np.testing.assert_array_equal(a[k], b[k], err_msg=k)


@pytest.mark.parametrize("whole_model", [True, False])
def test_h5_round_trip_in_both_keras_layouts(W, whole_model, tmp_path):
    p = tmp_path / "xception_v4_large_08_0.894.h5"
    blob = keras_io.export_keras_h5(W, str(p), whole_model=whole_model)
    assert len(blob) > 4 * weights.pack_blob(W).size            # every fp32 value is in there
    f = hdf5_min.File(str(p))
    root = f["model_weights"] if whole_model else f
    assert [n.decode() for n in root.attrs["layer_names"]] == ["xception", "dense_6", "dense_7"]
    assert root["xception"].attrs["weight_names"][0] == b"block1_conv1/kernel:0"
    assert root["xception/block2_sepconv1/depthwise_kernel:0"].shape == (3, 3, 64, 1)
    assert root["xception/conv2d_4/kernel:0"].shape == (1, 1, 64, 128)       # block2's residual projection, auto-named
    assert root["dense_7/dense_7/bias:0"].shape == (10,)
    _same(keras_io.load_keras_h5(str(p)), W)
    _same(keras_io.load_keras_h5(blob), W)
    # the blob the C ABI takes is identical
    np.testing.assert_array_equal(weights.pack_blob(keras_io.load_keras_h5(blob)), weights.pack_blob(W))


def test_npz_export_of_model_variables(W, tmp_path):
    named = {}
    for layer, vs in keras_io.keras_layout(W).items():
        named.update(vs)                                         # {v.name: v.numpy() for v in model.variables}
    p = tmp_path / "w.npz"
    np.savez(p, **named)
    _same(keras_io.load_keras_npz(p), W)


def test_auto_numbering_and_head_variants():
    W0 = weights.make_weights(5, "keras_default", head_inner=0)              # Dense(10) directly on the pooled features
    assert "dense_inner/kernel" not in W0
    # a fresh session: conv2d, conv2d_1, ... and a single dense layer called 'dense'
    lay = keras_io.keras_layout(W0, first_auto=0, first_dense=0)
    named = {f"model_weights/{layer}/{k}": v for layer, vs in lay.items() for k, v in vs.items()}
    assert "model_weights/xception/conv2d/kernel:0" in named and "model_weights/dense_0/dense_0/kernel:0" in named
    _same(keras_io.bundle_from_named_arrays(named), W0)
    # two-digit suffixes sort numerically (conv2d_9 before conv2d_10)
    lay = keras_io.keras_layout(W0, first_auto=8, first_dense=12)
    named = {f"{layer}/{k}": v for layer, vs in lay.items() for k, v in vs.items()}
    assert any(k.endswith("conv2d_10/kernel:0") for k in named)
    _same(keras_io.bundle_from_named_arrays(named), W0)
    # optimizer state is ignored
    named["optimizer_weights/training/Adam/dense_12/kernel/m:0"] = np.zeros((2048, 10), np.float32)
    _same(keras_io.bundle_from_named_arrays(named), W0)


def test_import_errors_name_the_variable(W):
    lay = keras_io.keras_layout(W)
    named = {f"{layer}/{k}": v for layer, vs in lay.items() for k, v in vs.items()}
    bad = dict(named)
    del bad["xception/block5_sepconv2_bn/moving_variance:0"]
    with pytest.raises(keras_io.WeightImportError, match="block5_sepconv2_bn/moving_variance"):
        keras_io.bundle_from_named_arrays(bad)
    bad = dict(named)
    bad["xception/block1_conv1/kernel:0"] = np.zeros((3, 3, 3, 16), np.float32)
    with pytest.raises(keras_io.WeightImportError, match=r"block1_conv1/kernel: shape \(3, 3, 3, 16\)"):
        keras_io.bundle_from_named_arrays(bad)
    bad = dict(named)
    bad["xception/block15_sepconv1/pointwise_kernel:0"] = np.zeros((1, 1, 8, 8), np.float32)
    with pytest.raises(keras_io.WeightImportError, match="no place for.*block15_sepconv1"):
        keras_io.bundle_from_named_arrays(bad)
    bad = {k: v for k, v in named.items() if "conv2d_7" not in k}
    with pytest.raises(keras_io.WeightImportError, match="4 unnamed residual convolutions"):
        keras_io.bundle_from_named_arrays(bad)
    bad = dict(named)
    bad["dense_9/dense_9/kernel:0"] = np.zeros((10, 10), np.float32)
    with pytest.raises(keras_io.WeightImportError, match="1 or 2 Dense layers"):
        keras_io.bundle_from_named_arrays(bad)


def test_spec_known_answers_survive_the_import(W):
    """21 067 390 parameters with the 100-unit head"""
    b = keras_io.load_keras_h5(keras_io.export_keras_h5(W))
    assert sum(a.size for a in b.values()) == 21_067_390
    assert len(b) == len(arch.weight_spec(100))
