"""hdf5_min: the reader against a file written by the real HDF5 library, writer -> reader round trips, a
hand-assembled chunked + deflate + shuffle dataset, and error reporting.  CPU only."""
import os
import zlib

import numpy as np
import pytest

import kdl_b200  # noqa: F401
from kdl_b200 import hdf5_min as h5


def _scipy_matlab_file():
    try:
        import scipy.io
    except ImportError:
        return None
    p = os.path.join(os.path.dirname(scipy.io.__file__), "matlab", "tests", "data", "testhdf5_7.4_GLNX86.mat")
    return p if os.path.exists(p) else None


@pytest.mark.skipif(_scipy_matlab_file() is None, reason="scipy's MATLAB 7.3 test file is not installed")
def test_reads_a_file_written_by_libhdf5():
    """MATLAB 7.4 wrote this with the HDF5 C library: 512-byte user block, superblock v0, symbol-table groups, a contiguous
    float64 dataset with a fixed-length string attribute.  Content per scipy's own test suite: arange(0, 2*pi, pi/4)."""
    f = h5.File(_scipy_matlab_file())
    assert f.keys() == ["testdouble"]
    n = f["testdouble"]
    assert n.is_dataset and n.attrs["MATLAB_class"] == b"double"
    a = n.read()
    assert a.dtype == np.float64 and a.shape == (9, 1)
    np.testing.assert_array_equal(a.ravel(), np.arange(0, 2 * np.pi + 1e-9, np.pi / 4))


def test_writer_reader_round_trip_nested_groups_dtypes_attributes():
    rng = np.random.default_rng(0)
    tree = {
        ".attrs": {"backend": np.bytes_(b"tensorflow"), "layer_names": np.array([b"a", b"bb", b"ccc"])},
        "f32": rng.standard_normal((3, 3, 4, 5)).astype(np.float32),
        "f64": rng.standard_normal((7,)),
        "i32": (np.arange(12, dtype=np.int32).reshape(3, 4), {"note": np.bytes_(b"ints"), "scale": np.float32(0.5)}),
        "u8": rng.integers(0, 256, (2, 5), dtype=np.uint8),
        "scalar": np.float32(3.25),
        "empty": np.zeros((0, 4), np.float32),
        "g": {".attrs": {"weight_names": np.array([b"g/x:0", b"g/sub/y:0"])},
              "x:0": np.ones((2, 2), np.float32),
              "sub": {"y:0": np.full((5,), 7, np.int64)},
              "hollow": {}},
    }
    blob = h5.write_file(None, tree)
    f = h5.File(blob)
    assert sorted(f.keys()) == sorted(k for k in tree if k != ".attrs")
    assert f.attrs["backend"] == b"tensorflow"
    assert list(f.attrs["layer_names"]) == [b"a", b"bb", b"ccc"]
    for k in ("f32", "f64", "u8", "empty"):
        got = f[k].read()
        assert got.dtype == tree[k].dtype and got.shape == tree[k].shape
        np.testing.assert_array_equal(got, tree[k])
    np.testing.assert_array_equal(f["i32"].read(), tree["i32"][0])
    assert f["i32"].attrs["note"] == b"ints" and f["i32"].attrs["scale"] == np.float32(0.5)
    assert f["scalar"].read() == np.float32(3.25) and f["scalar"].shape == ()
    assert list(f["g"].attrs["weight_names"]) == [b"g/x:0", b"g/sub/y:0"]
    np.testing.assert_array_equal(f["g/sub/y:0"].read(), np.full((5,), 7, np.int64))
    assert f["g/hollow"].is_group and f["g/hollow"].keys() == []
    assert "g/x:0" in f and "g/nope" not in f
    assert sorted(h5.read_tree(blob)) == ["/empty", "/f32", "/f64", "/g/sub/y:0", "/g/x:0", "/i32", "/scalar", "/u8"]


def test_groups_with_hundreds_of_members_round_trip(tmp_path):
    tree = {f"layer_{i:03d}": {"kernel:0": np.full((2, 3), i, np.float32)} for i in range(300)}
    p = tmp_path / "many.h5"
    h5.write_file(str(p), tree)
    f = h5.File(str(p))
    assert len(f.keys()) == 300
    for i in (0, 17, 299):
        np.testing.assert_array_equal(f[f"layer_{i:03d}/kernel:0"].read(), np.full((2, 3), i, np.float32))


def _chunked_file(a, chunk, deflate=True, shuffle=True):
    """assemble by hand what h5py writes for create_dataset(..., chunks=chunk, compression='gzip', shuffle=True)"""
    w = h5._W()
    w.put(bytes(96))
    es = a.dtype.itemsize
    entries = []
    for i0 in range(0, a.shape[0], chunk[0]):
        for j0 in range(0, a.shape[1], chunk[1]):
            c = np.zeros(chunk, a.dtype)
            part = a[i0:i0 + chunk[0], j0:j0 + chunk[1]]
            c[:part.shape[0], :part.shape[1]] = part
            raw = c.tobytes()
            if shuffle:
                raw = np.frombuffer(raw, np.uint8).reshape(-1, es).T.tobytes()
            if deflate:
                raw = zlib.compress(raw, 4)
            w.align()
            entries.append(((i0, j0), w.put(raw), len(raw)))
    w.align()
    node = bytearray(b"TREE" + bytes([1, 0]) + h5._u(len(entries), 2) + h5._u(h5.UNDEF, 8) + h5._u(h5.UNDEF, 8))
    for (i0, j0), addr, n in entries:
        node += h5._u(n, 4) + h5._u(0, 4) + h5._u(i0, 8) + h5._u(j0, 8) + h5._u(0, 8) + h5._u(addr, 8)
    node += h5._u(0, 4) + h5._u(0, 4) + h5._u(a.shape[0], 8) + h5._u(a.shape[1], 8) + h5._u(0, 8)     # final key
    bt = w.put(bytes(node))
    filters = []
    if shuffle:
        filters.append(h5._u(2, 2) + h5._u(0, 2) + h5._u(1, 2) + h5._u(1, 2) + h5._u(es, 4) + bytes(4))
    if deflate:
        filters.append(h5._u(1, 2) + h5._u(0, 2) + h5._u(1, 2) + h5._u(1, 2) + h5._u(4, 4) + bytes(4))
    msgs = [h5._msg(0x0001, h5._ds_msg(a.shape)), h5._msg(0x0003, h5._dt_msg(a.dtype)),
            h5._msg(0x0008, bytes([3, 2, 3]) + h5._u(bt, 8) + h5._u(chunk[0], 4) + h5._u(chunk[1], 4) + h5._u(es, 4))]
    if filters:
        msgs.append(h5._msg(0x000B, bytes([1, len(filters)]) + bytes(6) + b"".join(filters)))
    ds = h5._object_header(w, msgs)
    # root group holding the one dataset
    w.align()
    heap_data = w.put(bytes(8) + b"data\0\0\0\0")
    heap = w.put(b"HEAP" + bytes(4) + h5._u(16, 8) + h5._u(h5.UNDEF, 8) + h5._u(heap_data, 8))
    snod = w.put(b"SNOD" + bytes([1, 0]) + h5._u(1, 2) + h5._u(8, 8) + h5._u(ds, 8) + bytes(24))
    tree = w.put(b"TREE" + bytes([0, 0]) + h5._u(1, 2) + h5._u(h5.UNDEF, 8) + h5._u(h5.UNDEF, 8) + h5._u(0, 8) + h5._u(snod, 8) + h5._u(8, 8))
    root = h5._object_header(w, [h5._msg(0x0011, h5._u(tree, 8) + h5._u(heap, 8))])
    w.align()
    w.patch(0, h5.SIGNATURE + bytes([0, 0, 0, 0, 0, 8, 8, 0]) + h5._u(4, 2) + h5._u(16, 2) + h5._u(0, 4) + h5._u(0, 8) + h5._u(h5.UNDEF, 8)
            + h5._u(w.tell(), 8) + h5._u(h5.UNDEF, 8) + h5._u(0, 8) + h5._u(root, 8) + h5._u(1, 4) + bytes(4) + h5._u(tree, 8) + h5._u(heap, 8))
    return bytes(w.buf)


@pytest.mark.parametrize("deflate,shuffle", [(False, False), (True, False), (True, True)])
def test_chunked_datasets_with_deflate_and_shuffle(deflate, shuffle):
    a = np.random.default_rng(3).standard_normal((10, 7)).astype(np.float32)
    got = h5.File(_chunked_file(a, (4, 3), deflate, shuffle))["data"].read()     # ragged edge chunks in both dimensions
    assert got.dtype == np.float32
    np.testing.assert_array_equal(got, a)


def test_errors_name_the_problem():
    with pytest.raises(h5.Hdf5Error, match="not an HDF5 file"):
        h5.File(b"PK\x03\x04" + bytes(2000))
    blob = bytearray(h5.write_file(None, {"x": np.zeros(100, np.float32)}))
    with pytest.raises(h5.Hdf5Error, match="truncated"):
        h5.File(bytes(blob[:300]))
    f = h5.File(bytes(blob))
    with pytest.raises(KeyError):
        f["y"]
    with pytest.raises(h5.Hdf5Error, match="not a dataset"):
        f.read()
    v2 = bytearray(blob)
    v2[8] = 7
    with pytest.raises(h5.Hdf5Error, match="superblock version 7"):
        h5.File(bytes(v2))
    with pytest.raises(h5.Hdf5Error, match="cannot write dtype"):
        h5.write_file(None, {"c": np.zeros(3, np.complex64)})


def test_random_trees_round_trip():
    """property test: any nested tree of float / int arrays with string attributes survives write -> read"""
    hyp = pytest.importorskip("hypothesis")
    st = pytest.importorskip("hypothesis.strategies")
    hnp = pytest.importorskip("hypothesis.extra.numpy")
    names = st.text(alphabet="abcdefghijklmnopqrstuvwxyz_0123456789:", min_size=1, max_size=12)
    arrays = st.one_of(
        hnp.arrays(np.float32, hnp.array_shapes(min_dims=0, max_dims=4, max_side=5), elements=st.floats(-1e6, 1e6, width=32)),
        hnp.arrays(np.int64, hnp.array_shapes(min_dims=1, max_dims=2, max_side=6), elements=st.integers(-2**40, 2**40)),
        hnp.arrays(np.uint8, hnp.array_shapes(min_dims=1, max_dims=3, max_side=4)))
    trees = st.recursive(st.dictionaries(names, arrays, max_size=4),
                         lambda children: st.dictionaries(names, st.one_of(arrays, children), max_size=4), max_leaves=12)

    def flatten(t, prefix=""):
        out = {}
        for k, v in t.items():
            if isinstance(v, dict):
                out.update(flatten(v, prefix + "/" + k))
            else:
                out[prefix + "/" + k] = v
        return out

    @hyp.settings(max_examples=60, deadline=None)
    @hyp.given(trees, st.lists(st.binary(min_size=1, max_size=9).filter(lambda b: b"\0" not in b), max_size=3))
    def check(tree, strs):
        if strs:
            tree = dict(tree)
            tree[".attrs"] = {"names": np.array(strs)}
        blob = h5.write_file(None, tree)
        f = h5.File(blob)
        got = {k: n.read() for k, n in f.walk()}
        want = flatten({k: v for k, v in tree.items() if k != ".attrs"})
        assert sorted(got) == sorted(want)
        for k in want:
            assert got[k].dtype == want[k].dtype and got[k].shape == want[k].shape
            np.testing.assert_array_equal(got[k], want[k])
        if strs:
            assert [bytes(x) for x in f.attrs["names"]] == strs

    check()
