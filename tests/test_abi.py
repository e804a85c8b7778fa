"""CPU-only checks of the boundary: the library builds/loads without a GPU and exports every symbol that
include/kdl.h declares; host-side argument validation that needs no device."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import kdl_b200  # noqa: F401
    from kdl_b200 import _lib, build
    if not os.path.exists(_lib.LIB_PATH):
        build.build_lib()
    return _lib.load()


def test_header_symbols_all_exported(lib):
    from kdl_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "kdl.h")).read()
    declared = set(re.findall(r"\b(kdl_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)
    for s in declared:
        assert hasattr(lib, s), s


def test_weight_count_matches_spec(lib):
    from kdl_b200 import arch
    assert lib.kdl_weight_count(100) == arch.param_count(100) == 21_067_390
    assert lib.kdl_weight_count(0) == arch.param_count(0)
    assert lib.kdl_weight_count(37) == arch.param_count(37)


def test_version_string_and_null_arguments(lib):
    from kdl_b200 import _lib
    assert b"sm_100a" in lib.kdl_version()
    assert lib.kdl_create(None, None) == _lib.KDL_EINVAL
    assert b"null" in lib.kdl_last_error(None)
    assert lib.kdl_load_weights(None, None, 0) == _lib.KDL_EINVAL
    assert lib.kdl_destroy(None) == _lib.KDL_OK


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from kdl_b200 import engine, weights
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        engine.Model(weights.make_weights(1, "keras_default"))


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "kubernetes-deep-learning_b200")
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, fn)).read()
                assert "oracle" not in src.replace("no CPU fallback", ""), os.path.join(dp, fn)


def test_gateway_labels_match_reference_order():
    from kdl_b200 import gateway
    assert gateway.labels == ['dress', 'hat', 'longsleeve', 'outwear', 'pants', 'shirt', 'shoes', 'shorts',
                              'skirt', 't-shirt']
