"""A small pure-Python HDF5 reader (and a matching writer) for Keras weight files.

The clothing model ships as `xception_v4_large_08_0.894.h5` (convert.py:4, guide.md:176), a Keras HDF5 file; h5py is not installable
here, so this module reads the subset of the HDF5 file format that h5py/libhdf5 write for such files with the default
("earliest") format settings, following the published HDF5 File Format Specification 2.0/3.0:

  superblock v0/v1 (also found behind a 512·2^n-byte user block), old-style groups (symbol-table message -> v1 B-tree -> symbol
  table nodes -> local heap), version-1 object headers with continuation blocks, dataspace v1/v2, datatypes fixed-point / IEEE
  float / fixed-length string / variable-length string (global heap), data layout v1-v3 compact / contiguous / chunked (v1 chunk
  B-tree; deflate and shuffle filters), attribute messages v1-v3.

Not supported (raises Hdf5Error naming the feature): superblock v2/v3 files whose groups use link messages / fractal heaps
(h5py `libver='latest'`), compound / array / reference datatypes, external storage, virtual datasets.

`write_file` produces files in the same subset (contiguous datasets, fixed-length string attributes); it exists so that the
importer can be tested offline and so that a bundle can be exported in Keras layout.  The reader is pinned against a file written
by the real HDF5 library (scipy's MATLAB 7.3 test file) in tests/test_hdf5_min.py.
"""
import zlib
from typing import Dict, List, Optional, Tuple, Union

import numpy as np

SIGNATURE = b"\x89HDF\r\n\x1a\n"
UNDEF = 0xFFFFFFFFFFFFFFFF


class Hdf5Error(ValueError):
    pass


# ================================================================================================================================ reader
class _Datatype:
    def __init__(self, cls, size, np_dtype=None, vlen_string=False, strpad=None):
        self.cls, self.size, self.np_dtype, self.vlen_string, self.strpad = cls, size, np_dtype, vlen_string, strpad


class Node:
    """A group (children: dict name -> Node) or a dataset (`read()` -> ndarray), with `attrs`."""

    def __init__(self, f, addr, name):
        self._f, self._addr, self.name = f, addr, name
        self.attrs: Dict[str, object] = {}
        self._children: Optional[Dict[str, "Node"]] = None
        self._stab = None
        self._layout = self._dtype = self._shape = None
        self._filters: List[Tuple[int, Tuple[int, ...]]] = []

    @property
    def is_group(self):
        return self._stab is not None

    @property
    def is_dataset(self):
        return self._layout is not None

    @property
    def shape(self):
        return self._shape

    def keys(self):
        return list(self.children)

    @property
    def children(self) -> Dict[str, "Node"]:
        if self._children is None:
            self._children = {}
            if self._stab is not None:
                for nm, addr in self._f._iter_group(*self._stab):
                    self._children[nm] = self._f._node(addr, (self.name.rstrip("/") + "/" + nm))
        return self._children

    def __getitem__(self, path: str) -> "Node":
        n = self
        for part in [p for p in path.split("/") if p]:
            if part not in n.children:
                raise KeyError(f"{n.name or '/'} has no member {part!r}")
            n = n.children[part]
        return n

    def __contains__(self, path):
        try:
            self[path]
            return True
        except KeyError:
            return False

    def read(self) -> np.ndarray:
        if not self.is_dataset:
            raise Hdf5Error(f"{self.name} is not a dataset")
        return self._f._read_dataset(self)

    def walk(self):
        """yields (path, Node) for every dataset below this node"""
        if self.is_dataset:
            yield self.name, self
        elif self.is_group:
            for c in self.children.values():
                yield from c.walk()


class File(Node):
    def __init__(self, path_or_bytes: Union[str, bytes]):
        if isinstance(path_or_bytes, (bytes, bytearray, memoryview)):
            self._b = bytes(path_or_bytes)
        else:
            with open(path_or_bytes, "rb") as fh:
                self._b = fh.read()
        self._cache: Dict[int, Node] = {}
        base = 0
        while True:     # the superblock sits at 0, 512, 1024, 2048, ... (after an optional user block)
            if self._b[base:base + 8] == SIGNATURE:
                break
            base = 512 if base == 0 else base * 2
            if base + 8 > len(self._b):
                raise Hdf5Error("not an HDF5 file (signature not found)")
        b = self._b
        ver = b[base + 8]
        if ver in (0, 1):
            self.O, self.L = b[base + 13], b[base + 14]
            p = base + 24 + (4 if ver == 1 else 0)
            self.base = self._uint(p, self.O)      # every address in the file is relative to this (MATLAB: 512, h5py: 0)
            self._sb_off = base
            p += 4 * self.O             # base, free-space info, end of file, driver info
            # root group symbol table entry
            hdr = self._uint(p + self.O, self.O)
            root_addr = hdr
        elif ver in (2, 3):
            self.O, self.L = b[base + 9], b[base + 10]
            self.base = self._uint(base + 12, self.O)
            self._sb_off = base
            root_addr = self._uint(base + 12 + 3 * self.O, self.O)
        else:
            raise Hdf5Error(f"superblock version {ver} is not supported")
        if self.O != 8 or self.L != 8:
            raise Hdf5Error(f"only 8-byte offsets/lengths are supported (file has {self.O}/{self.L})")
        Node.__init__(self, self, root_addr, "")
        self._parse_object(self)

    # ---------------------------------------------------------------------------------------------------------- primitives
    def _uint(self, off, n):
        return int.from_bytes(self._b[off:off + n], "little")

    def _abs(self, addr):
        return addr + self.base

    def _node(self, addr, name) -> Node:
        if addr in self._cache:
            return self._cache[addr]
        n = Node(self, addr, name)
        self._cache[addr] = n
        self._parse_object(n)
        return n

    # ---------------------------------------------------------------------------------------------------------- object headers
    def _messages(self, addr):
        b = self._b
        p = self._abs(addr)
        if p + 16 > len(b):
            raise Hdf5Error(f"file is truncated or corrupt: object header at {addr:#x} lies beyond the end of the file")
        if b[p:p + 4] == b"OHDR":
            raise Hdf5Error("version-2 object headers (h5py libver='latest') are not supported; re-save with the default libver")
        if b[p] != 1:
            raise Hdf5Error(f"object header version {b[p]} at {addr:#x} is not supported")
        nmsg = self._uint(p + 2, 2)
        size = self._uint(p + 8, 4)
        blocks = [(p + 16, size)]
        out = []
        while blocks and len(out) < nmsg:
            q, left = blocks.pop(0)
            end = q + left
            while q + 8 <= end and len(out) < nmsg:
                mtype, msize, flags = self._uint(q, 2), self._uint(q + 2, 2), b[q + 4]
                data = b[q + 8:q + 8 + msize]
                q += 8 + msize
                if mtype == 0x0010:
                    blocks.append((self._abs(int.from_bytes(data[:8], "little")), int.from_bytes(data[8:16], "little")))
                if flags & 0x02:
                    raise Hdf5Error("shared header messages are not supported")
                out.append((mtype, data))
        return out

    def _parse_object(self, n: Node):
        for mtype, d in self._messages(n._addr):
            if mtype == 0x0011:                                   # symbol table: B-tree address, local heap address
                n._stab = (int.from_bytes(d[:8], "little"), int.from_bytes(d[8:16], "little"))
            elif mtype == 0x0001:
                n._shape = self._dataspace(d)
            elif mtype == 0x0003:
                n._dtype = self._datatype(d)[0]
            elif mtype == 0x0008:
                n._layout = self._layout_msg(d)
            elif mtype == 0x000B:
                n._filters = self._filters_msg(d)
            elif mtype == 0x000C:
                k, v = self._attribute(d)
                n.attrs[k] = v
            elif mtype in (0x0002, 0x0006):
                raise Hdf5Error(f"{n.name}: new-style groups (link messages) are not supported; re-save with the default libver")
            elif mtype == 0x0007:
                raise Hdf5Error(f"{n.name}: external data files are not supported")

    # ---------------------------------------------------------------------------------------------------------- message bodies
    @staticmethod
    def _dataspace(d) -> Optional[Tuple[int, ...]]:
        ver, rank, flags = d[0], d[1], d[2]
        if ver == 1:
            p = 8
        elif ver == 2:
            if d[3] == 2:
                return None       # null dataspace
            p = 4
        else:
            raise Hdf5Error(f"dataspace version {ver}")
        return tuple(int.from_bytes(d[p + 8 * i:p + 8 * i + 8], "little") for i in range(rank))

    def _datatype(self, d) -> Tuple[_Datatype, int]:
        cls, ver = d[0] & 0x0F, d[0] >> 4
        bits = d[1] | (d[2] << 8) | (d[3] << 16)
        size = int.from_bytes(d[4:8], "little")
        if cls == 0:
            order = ">" if bits & 1 else "<"
            signed = "i" if bits & 8 else "u"
            return _Datatype(cls, size, np.dtype(f"{order}{signed}{size}")), 8 + 4
        if cls == 1:
            order = ">" if bits & 1 else "<"
            if size not in (2, 4, 8):
                raise Hdf5Error(f"{size}-byte floating-point type")
            return _Datatype(cls, size, np.dtype(f"{order}f{size}")), 8 + 12
        if cls == 3:
            return _Datatype(cls, size, np.dtype(f"S{size}"), strpad=bits & 0x0F), 8
        if cls == 9:
            base, used = self._datatype(d[8:])
            if (bits & 0x0F) != 1:
                raise Hdf5Error("variable-length sequences are not supported (only variable-length strings)")
            return _Datatype(cls, size, None, vlen_string=True), 8 + used
        names = {2: "time", 4: "bitfield", 5: "opaque", 6: "compound", 7: "reference", 8: "enum", 10: "array"}
        raise Hdf5Error(f"datatype class {names.get(cls, cls)} is not supported")

    def _layout_msg(self, d):
        ver = d[0]
        if ver == 3:
            lc = d[1]
            if lc == 0:
                n = int.from_bytes(d[2:4], "little")
                return ("compact", d[4:4 + n])
            if lc == 1:
                return ("contiguous", int.from_bytes(d[2:10], "little"), int.from_bytes(d[10:18], "little"))
            if lc == 2:
                nd = d[2]
                bt = int.from_bytes(d[3:11], "little")
                dims = tuple(int.from_bytes(d[11 + 4 * i:15 + 4 * i], "little") for i in range(nd))
                return ("chunked", bt, dims)
            raise Hdf5Error(f"data layout class {lc}")
        if ver in (1, 2):
            nd, lc = d[1], d[2]
            p = 8
            addr = None
            if lc != 0:
                addr = int.from_bytes(d[p:p + 8], "little")
                p += 8
            dims = tuple(int.from_bytes(d[p + 4 * i:p + 4 * i + 4], "little") for i in range(nd))
            p += 4 * nd
            if lc == 1:
                return ("contiguous", addr, None)
            if lc == 2:
                es = int.from_bytes(d[p:p + 4], "little")
                return ("chunked", addr, dims + (es,))
            n = int.from_bytes(d[p:p + 4], "little")
            return ("compact", d[p + 4:p + 4 + n])
        raise Hdf5Error(f"data layout version {ver} is not supported (virtual / v4 chunk indexes need libver='latest')")

    @staticmethod
    def _filters_msg(d):
        ver, nf = d[0], d[1]
        p = 8 if ver == 1 else 2
        out = []
        for _ in range(nf):
            fid = int.from_bytes(d[p:p + 2], "little")
            p += 2
            nlen = 0
            if ver == 1 or fid >= 256:
                nlen = int.from_bytes(d[p:p + 2], "little")
                p += 2
            p += 2    # flags
            ncd = int.from_bytes(d[p:p + 2], "little")
            p += 2
            p += (nlen + 7) // 8 * 8 if ver == 1 else nlen
            cd = tuple(int.from_bytes(d[p + 4 * i:p + 4 * i + 4], "little") for i in range(ncd))
            p += 4 * ncd
            if ver == 1 and ncd % 2:
                p += 4
            out.append((fid, cd))
        return out

    def _attribute(self, d):
        ver = d[0]
        nsz, tsz, ssz = (int.from_bytes(d[2 + 2 * i:4 + 2 * i], "little") for i in range(3))
        if ver == 1:
            p = 8
            pad = lambda x: (x + 7) // 8 * 8   # noqa: E731
        elif ver in (2, 3):
            p = 8 if ver == 2 else 9
            pad = lambda x: x                  # noqa: E731
        else:
            raise Hdf5Error(f"attribute message version {ver}")
        name = d[p:p + nsz].split(b"\0")[0].decode("utf-8")
        p += pad(nsz)
        dt, _ = self._datatype(d[p:p + tsz])
        p += pad(tsz)
        shape = self._dataspace(d[p:p + ssz])
        p += pad(ssz)
        if shape is None:
            return name, None
        return name, self._decode(dt, shape, d[p:])

    # ---------------------------------------------------------------------------------------------------------- data
    def _global_heap_object(self, addr, index) -> bytes:
        b = self._b
        p = self._abs(addr)
        if b[p:p + 4] != b"GCOL":
            raise Hdf5Error(f"global heap collection expected at {addr:#x}")
        size = self._uint(p + 8, 8)
        q, end = p + 16, p + size
        while q + 16 <= end:
            idx = self._uint(q, 2)
            n = self._uint(q + 8, 8)
            if idx == 0:
                break
            if idx == index:
                return b[q + 16:q + 16 + n]
            q += 16 + (n + 7) // 8 * 8
        raise Hdf5Error(f"global heap object {index} not found at {addr:#x}")

    def _decode(self, dt: _Datatype, shape, raw: bytes):
        n = int(np.prod(shape)) if shape else 1
        if dt.vlen_string:
            out = []
            for i in range(n):
                e = raw[16 * i:16 * i + 16]
                ln, addr, idx = int.from_bytes(e[:4], "little"), int.from_bytes(e[4:12], "little"), int.from_bytes(e[12:16], "little")
                out.append(self._global_heap_object(addr, idx)[:ln].decode("utf-8") if addr not in (0, UNDEF) and ln else "")
            return out[0] if not shape else np.array(out, dtype=object).reshape(shape)
        a = np.frombuffer(raw, dtype=dt.np_dtype, count=n)
        if dt.cls == 3:
            if dt.strpad == 2:
                a = np.char.rstrip(a, b" ")
            return a.reshape(shape).copy() if shape else bytes(a[0])
        a = a.astype(dt.np_dtype.newbyteorder("="), copy=True).reshape(shape)
        return a if shape else a[()]

    def _read_dataset(self, n: Node) -> np.ndarray:
        dt, shape, lay = n._dtype, n._shape, n._layout
        if dt is None or shape is None:
            raise Hdf5Error(f"{n.name}: dataset without datatype/dataspace (null dataspace?)")
        count = int(np.prod(shape)) if shape else 1
        nbytes = count * dt.size
        if lay[0] == "compact":
            raw = lay[1][:nbytes]
        elif lay[0] == "contiguous":
            if n._filters:
                raise Hdf5Error(f"{n.name}: filters on a contiguous dataset")
            if lay[1] == UNDEF:
                raw = bytes(nbytes)                 # never written: fill value 0
            else:
                p = self._abs(lay[1])
                raw = self._b[p:p + nbytes]
                if len(raw) != nbytes:
                    raise Hdf5Error(f"{n.name}: file is truncated ({len(raw)} of {nbytes} bytes)")
        else:
            raw = self._read_chunked(n, nbytes)
        return self._decode(dt, shape, raw)

    def _read_chunked(self, n: Node, nbytes) -> bytes:
        _, bt, cdims = n._layout
        shape, es = n._shape, n._dtype.size
        rank = len(shape)
        if len(cdims) != rank + 1 or cdims[-1] != es:
            raise Hdf5Error(f"{n.name}: chunk dimensions {cdims} do not fit a rank-{rank} dataset of {es}-byte elements")
        chunk = cdims[:-1]
        out = np.zeros(shape, dtype=np.dtype(f"V{es}"))
        if bt != UNDEF:
            for offs, fmask, addr, csize in self._iter_chunks(bt, rank):
                raw = self._b[self._abs(addr):self._abs(addr) + csize]
                for i, (fid, cd) in reversed(list(enumerate(n._filters))):
                    if fmask & (1 << i):
                        continue
                    if fid == 1:
                        raw = zlib.decompress(raw)
                    elif fid == 2:       # shuffle
                        k = cd[0] if cd else es
                        raw = np.frombuffer(raw, np.uint8).reshape(k, -1).T.tobytes()
                    elif fid == 3:       # fletcher32: drop the trailing checksum
                        raw = raw[:-4]
                    else:
                        raise Hdf5Error(f"{n.name}: filter id {fid} is not supported")
                c = np.frombuffer(raw, dtype=out.dtype, count=int(np.prod(chunk))).reshape(chunk)
                sl_out = tuple(slice(o, min(o + c_, s)) for o, c_, s in zip(offs, chunk, shape))
                sl_in = tuple(slice(0, s.stop - s.start) for s in sl_out)
                out[sl_out] = c[sl_in]
        return out.tobytes()

    def _iter_chunks(self, addr, rank):
        b = self._b
        p = self._abs(addr)
        if b[p:p + 4] != b"TREE" or b[p + 4] != 1:
            raise Hdf5Error(f"chunk B-tree node expected at {addr:#x}")
        level, used = b[p + 5], self._uint(p + 6, 2)
        q = p + 8 + 16
        ksz = 8 + 8 * (rank + 1)
        for _ in range(used):
            csize, fmask = self._uint(q, 4), self._uint(q + 4, 4)
            offs = tuple(self._uint(q + 8 + 8 * i, 8) for i in range(rank))
            child = self._uint(q + ksz, 8)
            q += ksz + 8
            if level == 0:
                yield offs, fmask, child, csize
            else:
                yield from self._iter_chunks(child, rank)

    # ---------------------------------------------------------------------------------------------------------- groups
    def _heap_string(self, heap_addr, off) -> str:
        b = self._b
        p = self._abs(heap_addr)
        if b[p:p + 4] != b"HEAP":
            raise Hdf5Error(f"local heap expected at {heap_addr:#x}")
        data = self._abs(self._uint(p + 8 + 16, 8))
        end = b.index(b"\0", data + off)
        return b[data + off:end].decode("utf-8")

    def _iter_group(self, btree, heap):
        b = self._b
        p = self._abs(btree)
        if b[p:p + 4] == b"TREE":
            if b[p + 4] != 0:
                raise Hdf5Error(f"group B-tree node expected at {btree:#x}")
            used = self._uint(p + 6, 2)
            q = p + 8 + 16 + 8       # skip key 0
            for _ in range(used):
                yield from self._iter_group(self._uint(q, 8), heap)
                q += 16
        elif b[p:p + 4] == b"SNOD":
            nsym = self._uint(p + 6, 2)
            q = p + 8
            for _ in range(nsym):
                yield self._heap_string(heap, self._uint(q, 8)), self._uint(q + 8, 8)
                q += 40
        else:
            raise Hdf5Error(f"TREE or SNOD expected at {btree:#x}")


def read_tree(path_or_bytes) -> Dict[str, np.ndarray]:
    """every dataset of the file as {'/group/.../name': ndarray}"""
    f = File(path_or_bytes)
    return {k: n.read() for k, n in f.walk()}


# ================================================================================================================================ writer
class _W:
    def __init__(self):
        self.buf = bytearray()

    def tell(self):
        return len(self.buf)

    def align(self, a=8):
        self.buf.extend(bytes(-len(self.buf) % a))

    def put(self, b):
        off = len(self.buf)
        self.buf.extend(b)
        return off

    def patch(self, off, b):
        self.buf[off:off + len(b)] = b


def _u(v, n):
    return int(v).to_bytes(n, "little")


def _pad8(b):
    return b + bytes(-len(b) % 8)


def _dt_msg(dt: np.dtype) -> bytes:
    if dt.kind == "S":
        return bytes([0x13, 0x01, 0, 0]) + _u(dt.itemsize, 4)                       # class 3 v1, null-padded ASCII
    if dt.kind == "f":
        if dt.itemsize == 4:
            return bytes([0x11, 0x20, 31, 0]) + _u(4, 4) + _u(0, 2) + _u(32, 2) + bytes([23, 8, 0, 23]) + _u(127, 4)
        if dt.itemsize == 8:
            return bytes([0x11, 0x20, 63, 0]) + _u(8, 4) + _u(0, 2) + _u(64, 2) + bytes([52, 11, 0, 52]) + _u(1023, 4)
    if dt.kind in "iu":
        return bytes([0x10, 0x08 if dt.kind == "i" else 0, 0, 0]) + _u(dt.itemsize, 4) + _u(0, 2) + _u(8 * dt.itemsize, 2)
    raise Hdf5Error(f"cannot write dtype {dt}")


def _ds_msg(shape) -> bytes:
    return bytes([1, len(shape), 0, 0, 0, 0, 0, 0]) + b"".join(_u(s, 8) for s in shape)


def _msg(mtype, body) -> bytes:
    body = _pad8(body)
    if len(body) > 0xFFFF:
        raise Hdf5Error("header message larger than 64 KiB (split the attribute, as Keras does)")
    return _u(mtype, 2) + _u(len(body), 2) + bytes(4) + body


def _attr_msg(name: str, value) -> bytes:
    a = np.asarray(value)
    if a.dtype.kind == "U":
        a = np.char.encode(a, "utf-8")
    if a.dtype.kind == "S" and a.dtype.itemsize == 0:
        a = a.astype("S1")
    a = np.array(a, order="C")
    if a.dtype.byteorder == ">":
        a = a.astype(a.dtype.newbyteorder("<"))
    nm = name.encode("utf-8") + b"\0"
    dt, ds = _dt_msg(a.dtype), _ds_msg(a.shape)
    return _msg(0x000C, bytes([1, 0]) + _u(len(nm), 2) + _u(len(dt), 2) + _u(len(ds), 2) + _pad8(nm) + _pad8(dt) + _pad8(ds) + a.tobytes())


def _object_header(w: _W, msgs: List[bytes]) -> int:
    w.align()
    body = b"".join(msgs)
    return w.put(bytes([1, 0]) + _u(len(msgs), 2) + _u(1, 4) + _u(len(body), 4) + bytes(4) + body)


def _write_dataset(w: _W, arr: np.ndarray, attrs) -> int:
    a = np.array(arr, order="C")          # (ascontiguousarray would turn a scalar into shape (1,))
    if a.dtype.byteorder == ">":
        a = a.astype(a.dtype.newbyteorder("<"))
    w.align()
    data = w.put(a.tobytes()) if a.size else UNDEF
    msgs = [_msg(0x0001, _ds_msg(a.shape)), _msg(0x0003, _dt_msg(a.dtype)),
            _msg(0x0008, bytes([3, 1]) + _u(data, 8) + _u(a.nbytes, 8))]
    msgs += [_attr_msg(k, v) for k, v in (attrs or {}).items()]
    return _object_header(w, msgs)


def _write_group(w: _W, members: dict, attrs) -> Tuple[int, int, int]:
    """-> (object header address, B-tree address, heap address).  members: name -> dict (group) | ndarray | (ndarray, attrs);
    the reserved key '.attrs' of a group dict holds that group's attributes."""
    entries = []
    for name in sorted(k for k in members if k != ".attrs"):
        v = members[name]
        if isinstance(v, dict):
            oh, bt, hp = _write_group(w, v, v.get(".attrs"))
            entries.append((name, oh, 1, bt, hp))
        else:
            arr, at = v if isinstance(v, tuple) else (v, None)
            entries.append((name, _write_dataset(w, arr, at), 0, 0, 0))
    # local heap: offset 0 holds the empty string (B-tree key 0)
    heap = bytearray(b"\0" * 8)
    offs = []
    for name, *_ in entries:
        offs.append(len(heap))
        heap.extend(_pad8(name.encode("utf-8") + b"\0"))
    w.align()
    heap_data = w.put(bytes(heap))
    w.align()
    heap_addr = w.put(b"HEAP" + bytes([0, 0, 0, 0]) + _u(len(heap), 8) + _u(UNDEF, 8) + _u(heap_data, 8))
    # one symbol table node holding every member (the superblock declares a leaf K large enough), under a one-entry B-tree
    w.align()
    snod = bytearray(b"SNOD" + bytes([1, 0]) + _u(len(entries), 2))
    for (name, oh, cache, bt, hp), off in zip(entries, offs):
        snod += _u(off, 8) + _u(oh, 8) + _u(cache, 4) + bytes(4) + (_u(bt, 8) + _u(hp, 8) if cache else bytes(16))
    snod_addr = w.put(bytes(snod))
    w.align()
    if entries:
        bt_addr = w.put(b"TREE" + bytes([0, 0]) + _u(1, 2) + _u(UNDEF, 8) + _u(UNDEF, 8) + _u(0, 8) + _u(snod_addr, 8) + _u(offs[-1], 8))
    else:
        bt_addr = w.put(b"TREE" + bytes([0, 0]) + _u(0, 2) + _u(UNDEF, 8) + _u(UNDEF, 8) + _u(0, 8))
    msgs = [_msg(0x0011, _u(bt_addr, 8) + _u(heap_addr, 8))] + [_attr_msg(k, v) for k, v in (attrs or {}).items()]
    return _object_header(w, msgs), bt_addr, heap_addr


def write_file(path: Optional[str], tree: dict) -> bytes:
    """tree: nested dict; leaves are ndarrays or (ndarray, attrs); a group's attributes live under its '.attrs' key."""
    def widest(t):
        m = len([k for k in t if k != ".attrs"])
        return max([m] + [widest(v) for v in t.values() if isinstance(v, dict)])
    leaf_k = max(4, (widest(tree) + 1) // 2)
    if leaf_k > 0x7FFF:
        raise Hdf5Error("too many members in one group")
    w = _W()
    w.put(bytes(96))                       # superblock, patched below
    oh, bt, hp = _write_group(w, tree, tree.get(".attrs"))
    w.align()
    eof = w.tell()
    sb = (SIGNATURE + bytes([0, 0, 0, 0, 0, 8, 8, 0]) + _u(leaf_k, 2) + _u(16, 2) + _u(0, 4)
          + _u(0, 8) + _u(UNDEF, 8) + _u(eof, 8) + _u(UNDEF, 8)
          + _u(0, 8) + _u(oh, 8) + _u(1, 4) + bytes(4) + _u(bt, 8) + _u(hp, 8))
    assert len(sb) == 96
    w.patch(0, sb)
    out = bytes(w.buf)
    if path is not None:
        with open(path, "wb") as fh:
            fh.write(out)
    return out
