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


def test_ctypes_structs_mirror_the_header_field_for_field():
    """KdlConfig / KdlTuning / KdlProfileRow in _lib.py must list the same fields, in the same order and with the same array
    lengths, as include/kdl.h -- a silent mismatch would shift every kernel-selection switch by one slot."""
    import re
    import ctypes
    import kdl_b200  # noqa: F401
    from kdl_b200 import _lib
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "kdl.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    for name, cls in (("KdlConfig", _lib.KdlConfig), ("KdlTuning", _lib.KdlTuning), ("KdlProfileRow", _lib.KdlProfileRow)):
        body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (name, name), hdr, flags=re.S).group(1)
        want = []
        for decl in body.split(";"):
            m = re.match(r"\s*(\w+)\s+(\w+)(?:\[(\d+)\])?\s*$", decl)
            if m:
                want.append((m.group(2), m.group(1), int(m.group(3)) if m.group(3) else 0))
        got = []
        for fname, ftype in cls._fields_:
            n = ftype._length_ if hasattr(ftype, "_length_") else 0
            base = ftype._type_ if n else ftype
            cname = {ctypes.c_int32: "int32_t", ctypes.c_float: "float", ctypes.c_double: "double", ctypes.c_char: "char"}[base]
            got.append((fname, cname, n))
        assert got == want, (name, got, want)
