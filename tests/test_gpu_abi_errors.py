"""Error behaviour of the C ABI (include/kdl.h): every misuse returns a negative KDL_E* code with a message from
kdl_last_error, nothing aborts, and the model stays usable afterwards."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _cfg(_lib, **kw):
    d = dict(device=0, max_batch=2, input_size=96, head_inner=100, precision=_lib.PRECISION_BF16, gemm_impl=_lib.GEMM_AUTO,
             keep_taps=0, use_graphs=0)
    d.update(kw)
    return _lib.KdlConfig(**d)


def test_create_rejects_bad_configs():
    import kdl_b200  # noqa: F401
    from kdl_b200 import _lib
    lib = _lib.load()
    torch.cuda.init()
    for kw, frag in (({"max_batch": 0}, b"bad config"), ({"input_size": 8}, b"bad config"), ({"head_inner": 4096}, b"bad config"),
                     ({"precision": 7}, b"bad config"), ({"device": 99}, b"not present")):
        h = ctypes.c_void_p()
        cfg = _cfg(_lib, **kw)
        rc = lib.kdl_create(ctypes.byref(cfg), ctypes.byref(h))
        assert rc == _lib.KDL_EINVAL, kw
        assert frag in lib.kdl_last_error(None), (kw, lib.kdl_last_error(None))
        assert not h.value


def test_call_order_sizes_and_null_pointers_then_a_normal_forward():
    import kdl_b200  # noqa: F401
    from kdl_b200 import _lib, weights
    lib = _lib.load()
    torch.cuda.init()
    h = ctypes.c_void_p()
    cfg = _cfg(_lib)
    assert lib.kdl_create(ctypes.byref(cfg), ctypes.byref(h)) == _lib.KDL_OK
    try:
        x = torch.zeros((2, 96, 96, 3), dtype=torch.float32, device="cuda")
        y = torch.zeros((2, 10), dtype=torch.float32, device="cuda")
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        # forward before the weights are there
        assert lib.kdl_forward(h, x.data_ptr(), 2, y.data_ptr(), st) == _lib.KDL_ESTATE
        assert b"kdl_load_weights" in lib.kdl_last_error(h)
        # wrong number of weights
        W = weights.make_weights(1, "calibrated")
        blob = weights.pack_blob(W)
        assert lib.kdl_load_weights(h, blob.ctypes.data, blob.size - 1) == _lib.KDL_EINVAL
        assert lib.kdl_load_weights(h, None, blob.size) == _lib.KDL_EINVAL
        assert lib.kdl_load_weights(h, blob.ctypes.data, blob.size) == _lib.KDL_OK
        # batch limits and null buffers
        assert lib.kdl_forward(h, x.data_ptr(), 3, y.data_ptr(), st) == _lib.KDL_EINVAL
        assert b"max_batch" in lib.kdl_last_error(h)
        assert lib.kdl_forward(h, x.data_ptr(), 0, y.data_ptr(), st) == _lib.KDL_EINVAL
        assert lib.kdl_forward(h, None, 2, y.data_ptr(), st) == _lib.KDL_EINVAL
        assert lib.kdl_forward(h, x.data_ptr(), 2, None, st) == _lib.KDL_EINVAL
        hx = np.zeros((2, 96, 96, 3), np.float32)
        hy = np.zeros((2, 10), np.float32)
        assert lib.kdl_forward_host(h, None, 2, hy.ctypes.data) == _lib.KDL_EINVAL
        assert lib.kdl_predict_u8_host(h, None, 10, None, None, None, None, 1, hy.ctypes.data) == _lib.KDL_EINVAL
        assert lib.kdl_wait(h) in (_lib.KDL_OK, _lib.KDL_ESTATE)          # nothing in flight: must not hang
        # taps need keep_taps
        shape = (ctypes.c_int32 * 4)()
        relu = ctypes.c_int32()
        rc = lib.kdl_layer_output(h, b"block1_conv1_act", x.data_ptr(), x.numel(), ctypes.byref(shape), ctypes.byref(relu), st)
        assert rc in (_lib.KDL_EINVAL, _lib.KDL_ESTATE)
        assert lib.kdl_last_error(h)
        # after all of that the model still works, and a second identical call gives the same answer
        assert lib.kdl_forward_host(h, hx.ctypes.data, 2, hy.ctypes.data) == _lib.KDL_OK
        first = hy.copy()
        assert np.isfinite(first).all() and np.abs(first).max() > 0
        assert lib.kdl_forward_host(h, hx.ctypes.data, 2, hy.ctypes.data) == _lib.KDL_OK
        np.testing.assert_array_equal(hy, first)
        assert lib.kdl_last_launch_count(h) > 50
    finally:
        assert lib.kdl_destroy(h) == _lib.KDL_OK
    assert lib.kdl_destroy(None) in (_lib.KDL_OK, _lib.KDL_EINVAL)


def test_standalone_gemm_entry_validates_its_arguments():
    import kdl_b200  # noqa: F401
    from kdl_b200 import _lib
    lib = _lib.load()
    a = torch.zeros((128, 64), dtype=torch.bfloat16, device="cuda")
    b = torch.zeros((64, 64), dtype=torch.bfloat16, device="cuda")
    bias = torch.zeros(64, device="cuda")
    out = torch.zeros((128, 64), dtype=torch.bfloat16, device="cuda")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    assert lib.kdl_gemm_bf16(None, b.data_ptr(), bias.data_ptr(), None, out.data_ptr(), 128, 64, 64, 0, st) == _lib.KDL_EINVAL
    assert lib.kdl_gemm_bf16(a.data_ptr(), b.data_ptr(), bias.data_ptr(), None, out.data_ptr(), 128, 60, 64, 0, st) == _lib.KDL_EINVAL
    assert b"multiples of 8" in lib.kdl_last_error(None)
    assert lib.kdl_gemm_bf16(a.data_ptr(), b.data_ptr(), bias.data_ptr(), None, out.data_ptr(), 128, 64, 64, 0, st) == _lib.KDL_OK
    torch.cuda.synchronize()


def test_host_entry_points_validate_geometry_before_queueing_any_copy():
    """kdl_submit_u8_host must refuse inconsistent image metadata the way kdl_predict_u8_host does --
    an image that extends past total_bytes would make k_resize read past the staging buffer, and an illegal address is sticky
    for the whole process.  Both entry points share one check; afterwards the model still works."""
    import kdl_b200  # noqa: F401
    from kdl_b200 import _lib, engine, weights
    lib = _lib.load()
    m = engine.Model(weights.make_weights(1, "keras_default"), max_batch=2, input_size=96, precision="bf16", use_graphs=False)
    try:
        img = np.full((40, 50, 3), 7, np.uint8)
        pk = engine.PackedImages([img, img], pinned=True)
        out = torch.empty((2, 10), dtype=torch.float32).pin_memory()

        def call(fn, total=None, **over):
            arr = {"offset": pk.offset.copy(), "h": pk.h.copy(), "w": pk.w.copy(), "pitch": pk.pitch.copy()}
            ptr = {}
            for k in arr:
                if k in over and over[k] is None:
                    ptr[k] = None
                else:
                    if k in over:
                        arr[k][1] = over[k]
                    ptr[k] = arr[k].ctypes.data
            rc = fn(m.h, pk.host.data_ptr(), pk.total_bytes if total is None else total, ptr["offset"], ptr["h"], ptr["w"], ptr["pitch"],
                    2, out.data_ptr())
            return rc, lib.kdl_last_error(m.h)

        for fn in (lib.kdl_submit_u8_host, lib.kdl_predict_u8_host):
            rc, msg = call(fn, h=4000)                         # 4000 rows of 150 bytes do not fit the packed buffer
            assert rc == _lib.KDL_EINVAL and b"extends past total_bytes" in msg, msg
            rc, msg = call(fn, offset=pk.total_bytes)          # starts at the very end
            assert rc == _lib.KDL_EINVAL and b"extends past total_bytes" in msg, msg
            rc, msg = call(fn, total=100)
            assert rc == _lib.KDL_EINVAL and b"extends past total_bytes" in msg, msg
            rc, msg = call(fn, w=0)
            assert rc == _lib.KDL_EINVAL and b"bad geometry" in msg, msg
            rc, msg = call(fn, pitch=10)                       # pitch smaller than a row
            assert rc == _lib.KDL_EINVAL and b"bad geometry" in msg, msg
            rc, msg = call(fn, w=None)                         # null metadata array (used to be dereferenced)
            assert rc == _lib.KDL_EINVAL and b"null metadata" in msg, msg
        assert lib.kdl_wait(m.h) == _lib.KDL_OK                # nothing was queued
        rc, _ = call(lib.kdl_submit_u8_host)
        assert rc == _lib.KDL_OK and lib.kdl_wait(m.h) == _lib.KDL_OK
        want = m.predict_packed_host(pk)
        np.testing.assert_array_equal(out.numpy(), want)
    finally:
        m.close()
