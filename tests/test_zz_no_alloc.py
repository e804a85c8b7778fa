"""kdl_preprocess_u8_static -- the resize+scale entry point for callers that cannot allocate device memory (a request
thread that must not touch the allocator, a plugin inside another framework's memory budget, code without a KdlModel).
Its index tables live in a fixed device array inside the library rather than in cudaMalloc'ed buffers.

The check: called through the C ABI, twice in a row on one stream, its output must equal the regular entry point's
(kdl_preprocess_u8) bit for bit -- preprocessing is integer/index work, so the tolerance is zero, as in
test_gpu_preprocess.py.  The second call uses a smaller batch and other sizes, so the array still holds the first call's
leftovers when it runs.  Kept in a file pytest collects last."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _call_static(lib, imgs, S):
    import torch
    from kdl_b200.engine import PackedImages
    pk = PackedImages(imgs, pinned=False)
    d_src = pk.host.cuda()
    out = torch.empty((pk.B, S, S, 3), dtype=torch.float32, device="cuda")
    err = ctypes.create_string_buffer(256)
    rc = lib.kdl_preprocess_u8_static(d_src.data_ptr(), *pk.meta_ptrs(), pk.B, S, out.data_ptr(),
                                      ctypes.c_void_p(torch.cuda.current_stream().cuda_stream), err, 256)
    return rc, out, err.value.decode()


def test_static_scratch_entry_matches_regular_entry_twice_in_a_row():
    import torch
    import kdl_b200  # noqa: F401
    from kdl_b200 import _lib, engine, weights
    lib = _lib.load()
    torch.cuda.set_device(0)
    rng = np.random.default_rng(21)
    first = [rng.integers(0, 256, (int(rng.integers(1, 900)), int(rng.integers(1, 900)), 3), dtype=np.uint8)
             for _ in range(_lib.STATIC_MAX_BATCH)]                       # a full-capacity batch
    second = [rng.integers(0, 256, s + (3,), dtype=np.uint8) for s in [(534, 400), (1, 1), (299, 299), (64, 2048), (7, 3)]]
    m = engine.Model(weights.make_weights(1, "keras_default"), max_batch=_lib.STATIC_MAX_BATCH, precision="f32")
    try:
        for imgs in (first, second):                                      # two calls in a row, shared scratch
            rc, got, msg = _call_static(lib, imgs, 299)
            assert rc == 0, msg
            want = m.preprocess_uint8(imgs)                               # existing entry point: kdl_preprocess_u8
            torch.cuda.synchronize()
            assert got.shape == want.shape
            assert torch.equal(got, want), "static-scratch output differs from kdl_preprocess_u8"
    finally:
        m.close()


def test_static_scratch_entry_refuses_what_does_not_fit():
    import torch
    import kdl_b200  # noqa: F401
    from kdl_b200 import _lib
    lib = _lib.load()
    torch.cuda.set_device(0)
    img = np.zeros((8, 8, 3), np.uint8)
    rc, _, msg = _call_static(lib, [img] * (_lib.STATIC_MAX_BATCH + 1), 299)
    assert rc == _lib.KDL_EINVAL and "KDL_STATIC_MAX_BATCH" in msg
    rc, _, msg = _call_static(lib, [img], 16)
    assert rc == 0, msg
    from kdl_b200.engine import PackedImages
    pk = PackedImages([img], pinned=False)
    d = pk.host.cuda()
    err = ctypes.create_string_buffer(256)
    out = torch.empty(4, device="cuda")
    rc = lib.kdl_preprocess_u8_static(d.data_ptr(), *pk.meta_ptrs(), 1, _lib.STATIC_MAX_SIZE + 1, out.data_ptr(), None, err, 256)
    assert rc == _lib.KDL_EINVAL and b"KDL_STATIC_MAX_SIZE" in err.value


def test_static_scratch_entry_allocates_nothing_and_refuses_misaligned_output():
    """The property in the entry point's name: device memory in use is the same before and after a call (cudaMemGetInfo, so
    allocations made behind torch's caching allocator would show too), including with torch's cache emptied first so that any
    allocation would have to go to the driver.  And an output pointer that is not 16-byte aligned is refused on the host
    (KDL_EINVAL) instead of faulting in the kernel."""
    import torch
    import kdl_b200  # noqa: F401
    from kdl_b200 import _lib
    from kdl_b200.engine import PackedImages
    lib = _lib.load()
    torch.cuda.set_device(0)
    rng = np.random.default_rng(5)
    imgs = [rng.integers(0, 256, (120, 77, 3), dtype=np.uint8), rng.integers(0, 256, (300, 401, 3), dtype=np.uint8)]
    rc, ref, msg = _call_static(lib, imgs, 150)              # warm: the per-device symbol lookup happens on the first call
    assert rc == 0, msg
    pk = PackedImages(imgs, pinned=False)
    d_src = pk.host.cuda()
    out = torch.empty((2, 150, 150, 3), dtype=torch.float32, device="cuda")
    err = ctypes.create_string_buffer(256)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        assert lib.kdl_preprocess_u8_static(d_src.data_ptr(), *pk.meta_ptrs(), 2, 150, out.data_ptr(), st, err, 256) == 0, err.value
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, f"device memory in use changed across the calls: {free0} -> {free1}"
    assert torch.equal(out, ref)
    # misaligned output: a float view starting 4 bytes into an allocation
    big = torch.empty(2 * 150 * 150 * 3 + 4, dtype=torch.float32, device="cuda")
    rc = lib.kdl_preprocess_u8_static(d_src.data_ptr(), *pk.meta_ptrs(), 2, 150, big.data_ptr() + 4, st, err, 256)
    assert rc == _lib.KDL_EINVAL and b"16-byte aligned" in err.value
    torch.cuda.synchronize()                                  # nothing was launched, the context is healthy
