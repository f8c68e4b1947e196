"""Minimal HDF5 reader/writer for Keras ``.h5`` model files (no libhdf5, no h5py).

Replaces the ``h5py`` dependency behind ``tf.keras.models.load_model`` / ``Model.save``
(reference call sites: python-scripts/AUTOENCODER-TensorFlow-IO-Kafka/cardata-v1.py:199-202,
cardata-v3.py:227,261; notebook ``load_model`` nb2.0:888).

Only the structures Keras 2.2.4-tf actually writes are understood:
superblock v0, version-1 object headers with continuation blocks, symbol-table groups
(B-tree v1 + local heap + SNOD), contiguous/compact little-endian datasets, fixed-length and
variable-length (global heap) string attributes.  The writer emits the same subset so that a
file produced by :func:`write_h5` round-trips through :func:`H5File` and is laid out like the
files Keras itself produces.
"""
from __future__ import annotations

import struct
from typing import Dict, List, Optional, Tuple

import numpy as np

_SIG = b"\x89HDF\r\n\x1a\n"
_UNDEF = 0xFFFFFFFFFFFFFFFF


class H5Error(ValueError):
    pass


# ----------------------------------------------------------------------------------------------
# reader
# ----------------------------------------------------------------------------------------------
class _Datatype:
    __slots__ = ("cls", "size", "np_dtype", "vlen_str", "raw_len")

    def __init__(self, cls, size, np_dtype, vlen_str, raw_len):
        self.cls, self.size, self.np_dtype, self.vlen_str, self.raw_len = cls, size, np_dtype, vlen_str, raw_len


def _parse_datatype(buf: bytes, off: int) -> _Datatype:
    b0 = buf[off]
    cls, ver = b0 & 0x0F, b0 >> 4
    bits0 = buf[off + 1]
    size = struct.unpack_from("<I", buf, off + 4)[0]
    if cls == 0:  # fixed point
        signed = bool(bits0 & 0x08)
        order = ">" if (bits0 & 1) else "<"
        dt = np.dtype(f"{order}{'i' if signed else 'u'}{size}")
        return _Datatype(cls, size, dt, False, 8 + 4)
    if cls == 1:  # floating point
        order = ">" if (bits0 & 1) else "<"
        dt = np.dtype(f"{order}f{size}")
        return _Datatype(cls, size, dt, False, 8 + 12)
    if cls == 3:  # fixed-length string
        return _Datatype(cls, size, np.dtype(f"S{size}"), False, 8)
    if cls == 9:  # variable length
        vtype = bits0 & 0x0F
        base = _parse_datatype(buf, off + 8)
        return _Datatype(cls, size, None, vtype == 1, 8 + base.raw_len)
    raise H5Error(f"unsupported HDF5 datatype class {cls} (version {ver})")


def _parse_dataspace(buf: bytes, off: int) -> Tuple[int, ...]:
    ver, rank, flags = buf[off], buf[off + 1], buf[off + 2]
    if ver == 1:
        p = off + 8
    elif ver == 2:
        p = off + 4
    else:
        raise H5Error(f"unsupported dataspace version {ver}")
    return tuple(struct.unpack_from("<Q", buf, p + 8 * i)[0] for i in range(rank))


class H5Object:
    """A group or dataset: attributes, children (groups) and array data (datasets)."""

    def __init__(self, f: "H5File", addr: int, name: str):
        self._f, self.addr, self.name = f, addr, name
        self.attrs: Dict[str, object] = {}
        self._symtab: Optional[Tuple[int, int]] = None
        self._shape = None
        self._dtype: Optional[_Datatype] = None
        self._layout = None
        self._children: Optional[Dict[str, "H5Object"]] = None
        self._parse_header()

    # -- object header (version 1) -------------------------------------------------------------
    def _parse_header(self):
        buf = self._f.buf
        a = self.addr
        if buf[a] != 1:
            raise H5Error(f"object header version {buf[a]} at {a} not supported (only v1)")
        nmsg = struct.unpack_from("<H", buf, a + 2)[0]
        hsize = struct.unpack_from("<I", buf, a + 8)[0]
        blocks = [(a + 16, hsize)]
        seen = 0
        while blocks and seen < nmsg:
            p, left = blocks.pop(0)
            end = p + left
            while p + 8 <= end and seen < nmsg:
                mtype, msize, _mflags = struct.unpack_from("<HHB", buf, p)
                body = p + 8
                seen += 1
                if mtype == 0x10:
                    coff, clen = struct.unpack_from("<QQ", buf, body)
                    blocks.append((coff, clen))
                elif mtype == 0x11:
                    self._symtab = struct.unpack_from("<QQ", buf, body)
                elif mtype == 0x01:
                    self._shape = _parse_dataspace(buf, body)
                elif mtype == 0x03:
                    self._dtype = _parse_datatype(buf, body)
                elif mtype == 0x08:
                    self._layout = self._parse_layout(body)
                elif mtype == 0x0C:
                    k, v = self._parse_attribute(body)
                    self.attrs[k] = v
                p = body + msize

    def _parse_layout(self, off):
        buf = self._f.buf
        ver, cls = buf[off], buf[off + 1]
        if ver != 3:
            raise H5Error(f"data layout version {ver} not supported")
        if cls == 1:
            addr, size = struct.unpack_from("<QQ", buf, off + 2)
            return ("contiguous", addr, size)
        if cls == 0:
            size = struct.unpack_from("<H", buf, off + 2)[0]
            return ("compact", off + 4, size)
        raise H5Error("chunked datasets are not supported (Keras writes contiguous ones)")

    def _parse_attribute(self, off):
        buf = self._f.buf
        ver = buf[off]
        nsz, tsz, ssz = struct.unpack_from("<HHH", buf, off + 2)
        if ver == 1:
            pad = lambda n: (n + 7) & ~7
            p = off + 8
        elif ver in (2, 3):
            pad = lambda n: n
            p = off + 8 + (1 if ver == 3 else 0)
        else:
            raise H5Error(f"attribute message version {ver}")
        name = buf[p:p + nsz].split(b"\0", 1)[0].decode("utf-8")
        p += pad(nsz)
        dt = _parse_datatype(buf, p)
        p += pad(tsz)
        shape = _parse_dataspace(buf, p) if ssz else ()
        p += pad(ssz)
        return name, self._f._read_values(dt, shape, p)

    # -- groups ---------------------------------------------------------------------------------
    @property
    def is_group(self):
        return self._symtab is not None

    def keys(self) -> List[str]:
        return list(self._load_children().keys())

    def _load_children(self):
        if self._children is None:
            self._children = {}
            if self._symtab is not None:
                btree, heap = self._symtab
                heap_data = self._f._local_heap(heap)
                for nm_off, oaddr in self._f._walk_btree(btree):
                    end = self._f.buf.index(b"\0", heap_data + nm_off)
                    nm = self._f.buf[heap_data + nm_off:end].decode("utf-8")
                    self._children[nm] = H5Object(self._f, oaddr, self.name.rstrip("/") + "/" + nm)
        return self._children

    def __contains__(self, key):
        try:
            self[key]
            return True
        except KeyError:
            return False

    def __getitem__(self, path: str) -> "H5Object":
        node = self
        for part in [p for p in path.split("/") if p]:
            ch = node._load_children()
            if part not in ch:
                raise KeyError(f"{path!r}: no member {part!r} in {node.name!r}")
            node = ch[part]
        return node

    # -- datasets -------------------------------------------------------------------------------
    @property
    def shape(self):
        return self._shape

    @property
    def file_offset(self):
        return None if self._layout is None else self._layout[1]

    def read(self) -> np.ndarray:
        if self._layout is None or self._dtype is None:
            raise H5Error(f"{self.name} is not a dataset")
        return self._f._read_values(self._dtype, self._shape or (), self._layout[1])

    def visit(self, fn, prefix=""):
        for k, v in self._load_children().items():
            p = f"{prefix}/{k}" if prefix else k
            fn(p, v)
            if v.is_group:
                v.visit(fn, p)


class H5File(H5Object):
    def __init__(self, path_or_bytes):
        if isinstance(path_or_bytes, (bytes, bytearray, memoryview)):
            self.buf = bytes(path_or_bytes)
        else:
            with open(path_or_bytes, "rb") as fh:
                self.buf = fh.read()
        buf = self.buf
        if buf[:8] != _SIG:
            raise H5Error("not an HDF5 file (bad signature)")
        if buf[8] != 0:
            raise H5Error(f"superblock version {buf[8]} not supported (Keras/h5py default is 0)")
        if buf[13] != 8 or buf[14] != 8:
            raise H5Error("only 8-byte offsets/lengths supported")
        root_addr = struct.unpack_from("<Q", buf, 24 + 32 + 8)[0]
        super().__init__(self, root_addr, "/")

    def _local_heap(self, addr) -> int:
        if self.buf[addr:addr + 4] != b"HEAP":
            raise H5Error("bad local heap signature")
        return struct.unpack_from("<Q", self.buf, addr + 24)[0]

    def _walk_btree(self, addr):
        buf = self.buf
        if buf[addr:addr + 4] == b"SNOD":
            n = struct.unpack_from("<H", buf, addr + 6)[0]
            for i in range(n):
                e = addr + 8 + 40 * i
                yield struct.unpack_from("<QQ", buf, e)
            return
        if buf[addr:addr + 4] != b"TREE":
            raise H5Error("bad B-tree signature")
        used = struct.unpack_from("<H", buf, addr + 6)[0]
        p = addr + 24
        for i in range(used):
            child = struct.unpack_from("<Q", buf, p + 8 + 16 * i)[0]
            yield from self._walk_btree(child)

    def _gheap_object(self, coll_addr, index) -> bytes:
        buf = self.buf
        if buf[coll_addr:coll_addr + 4] != b"GCOL":
            raise H5Error("bad global heap signature")
        csize = struct.unpack_from("<Q", buf, coll_addr + 8)[0]
        p, end = coll_addr + 16, coll_addr + csize
        while p + 16 <= end:
            idx, _rc = struct.unpack_from("<HH", buf, p)
            osz = struct.unpack_from("<Q", buf, p + 8)[0]
            if idx == index:
                return buf[p + 16:p + 16 + osz]
            if idx == 0:
                break
            p += 16 + ((osz + 7) & ~7)
        raise H5Error(f"global heap object {index} not found")

    def _read_values(self, dt: _Datatype, shape, off):
        n = int(np.prod(shape)) if shape else 1
        if dt.cls == 9:
            if not dt.vlen_str:
                raise H5Error("only variable-length *strings* supported")
            out = []
            for i in range(n):
                ln, caddr, idx = struct.unpack_from("<IQI", self.buf, off + 16 * i)
                out.append(self._gheap_object(caddr, idx)[:ln].decode("utf-8"))
            return out[0] if not shape else np.array(out, dtype=object).reshape(shape)
        arr = np.frombuffer(self.buf, dtype=dt.np_dtype, count=n, offset=off)
        if dt.cls == 3:
            vals = [x.split(b"\0", 1)[0].decode("utf-8") for x in arr.tolist()]
            return vals[0] if not shape else np.array(vals, dtype=object).reshape(shape)
        arr = arr.astype(dt.np_dtype.newbyteorder("="), copy=True)
        return arr.reshape(shape) if shape else arr.reshape(())


# ----------------------------------------------------------------------------------------------
# writer (same subset; enough for Keras ``save`` layout: groups, f32/i64 datasets, string attrs)
# ----------------------------------------------------------------------------------------------
class _W:
    def __init__(self):
        self.b = bytearray()

    def tell(self):
        return len(self.b)

    def align(self, n=8):
        while len(self.b) % n:
            self.b.append(0)

    def write(self, data: bytes) -> int:
        self.align()
        at = len(self.b)
        self.b += data
        return at

    def patch(self, at, data):
        self.b[at:at + len(data)] = data


def _dt_msg(arr_dtype: np.dtype) -> bytes:
    if arr_dtype.kind == "f":
        size = arr_dtype.itemsize
        if size == 4:
            return struct.pack("<BBBBI", 0x11, 0x20, 0x1F, 0x00, 4) + struct.pack("<HHBBBBI", 0, 32, 23, 8, 0, 23, 127)
        return struct.pack("<BBBBI", 0x11, 0x20, 0x3F, 0x00, 8) + struct.pack("<HHBBBBI", 0, 64, 52, 11, 0, 52, 1023)
    if arr_dtype.kind in "iu":
        size = arr_dtype.itemsize
        return struct.pack("<BBBBI", 0x10, 0x08 if arr_dtype.kind == "i" else 0, 0, 0, size) + struct.pack("<HH", 0, size * 8)
    if arr_dtype.kind == "S":
        return struct.pack("<BBBBI", 0x13, 0x00, 0, 0, arr_dtype.itemsize)
    raise H5Error(f"cannot write dtype {arr_dtype}")


def _ds_msg(shape) -> bytes:
    if shape == ():
        return struct.pack("<BBBBI", 1, 0, 0, 0, 0)
    return struct.pack("<BBBBI", 1, len(shape), 1, 0, 0) + b"".join(struct.pack("<Q", d) for d in shape) * 2


def _pad8(b: bytes) -> bytes:
    return b + b"\0" * (-len(b) % 8)


def _msg(mtype: int, body: bytes, flags=0) -> bytes:
    body = _pad8(body)
    return struct.pack("<HHBBBB", mtype, len(body), flags, 0, 0, 0) + body


class _Node:
    def __init__(self):
        self.children: Dict[str, "_Node"] = {}
        self.attrs: Dict[str, object] = {}
        self.data: Optional[np.ndarray] = None


def _attr_msg(w: _W, name: str, value, gheap: List[bytes]) -> bytes:
    nm = name.encode("utf-8") + b"\0"
    if isinstance(value, str):
        # variable-length string → descriptor into the single global heap collection
        gheap.append(value.encode("utf-8"))
        idx = len(gheap)
        dt = struct.pack("<BBBBI", 0x19, 0x01, 0x01, 0x00, 16) + _dt_msg(np.dtype("S1"))
        ds = _ds_msg(())
        data = ("VLEN", len(value.encode("utf-8")), idx)
    else:
        arr = np.asarray(value)
        if arr.dtype.kind in "UO":
            enc = [str(s).encode("utf-8") for s in arr.ravel().tolist()]
            ln = max([len(e) for e in enc] + [1])
            arr = np.array(enc, dtype=f"S{ln}").reshape(arr.shape)
        arr = np.asarray(arr, order="C")
        if arr.dtype.kind in "fiu":
            arr = arr.astype(arr.dtype.newbyteorder("<"))
        dt, ds, data = _dt_msg(arr.dtype), _ds_msg(arr.shape), arr.tobytes()
    head = struct.pack("<BBHHH", 1, 0, len(nm), len(dt), len(ds))
    return head, _pad8(nm) + _pad8(dt) + _pad8(ds), data


def write_h5(path: str, tree: Dict, root_attrs: Optional[Dict] = None) -> None:
    """Write nested ``{name: ndarray | dict}`` with per-group ``'@attrs'`` dicts to ``path``."""
    def build(d) -> _Node:
        n = _Node()
        for k, v in d.items():
            if k == "@attrs":
                n.attrs.update(v)
            elif isinstance(v, dict):
                n.children[k] = build(v)
            else:
                c = _Node()
                c.data = np.asarray(v)
                n.children[k] = c
        return n

    root = build(tree)
    if root_attrs:
        root.attrs.update(root_attrs)

    w = _W()
    w.b += b"\0" * 96  # superblock placeholder (root object header follows at 96, as in h5py files)
    gheap: List[bytes] = []
    vlen_fixups: List[Tuple[int, int, int]] = []  # (position of 16-byte descriptor, length, index)

    def emit(node: _Node) -> int:
        # children first so that their addresses are known
        child_addr = {k: emit(c) for k, c in sorted(node.children.items())}
        msgs = []
        if node.data is not None:
            arr = np.asarray(node.data, order="C")
            if arr.dtype.kind in "fiu":
                arr = arr.astype(arr.dtype.newbyteorder("<"))
            raw = arr.tobytes()
            daddr = w.write(raw) if raw else _UNDEF
            msgs.append(_msg(0x01, _ds_msg(arr.shape)))
            msgs.append(_msg(0x03, _dt_msg(arr.dtype), flags=1))
            msgs.append(_msg(0x05, struct.pack("<BBBB", 2, 2, 2, 0)))  # fill value v2: late alloc, never written
            msgs.append(_msg(0x08, struct.pack("<BBQQ", 3, 1, daddr, len(raw))))
        else:
            # local heap with child names, one SNOD, one B-tree node
            names = sorted(child_addr)
            heap = bytearray(b"\0" * 8)
            offs = {}
            for nm in names:
                offs[nm] = len(heap)
                heap += nm.encode("utf-8") + b"\0"
                while len(heap) % 8:
                    heap.append(0)
            heap_size = max(len(heap) + 16, 88)
            free_off = len(heap)
            heap += b"\0" * (heap_size - len(heap))
            struct.pack_into("<QQ", heap, free_off, 1, heap_size - free_off)  # free block: next=1 (none), size
            hdata = w.write(bytes(heap))
            haddr = w.write(b"HEAP" + struct.pack("<BBBBQQQ", 0, 0, 0, 0, heap_size, free_off, hdata))
            K = 4 * 2  # leaf K=4 → SNOD holds up to 8 entries; chain as many SNODs as needed
            groups = [names[i:i + K] for i in range(0, len(names), K)] or [[]]
            snods = []
            for g in groups:
                body = b"SNOD" + struct.pack("<BBH", 1, 0, len(g))
                for nm in g:
                    body += struct.pack("<QQII", offs[nm], child_addr[nm], 0, 0) + b"\0" * 16
                body += b"\0" * (8 + 40 * K - len(body))
                snods.append(w.write(body))
            if len(groups) > 2 * 16:
                raise H5Error("group too large for a single B-tree node")
            tree_b = b"TREE" + struct.pack("<BBHQQ", 0, 0, len(groups), _UNDEF, _UNDEF)
            tree_b += struct.pack("<Q", 0)
            for g, s in zip(groups, snods):
                tree_b += struct.pack("<QQ", s, offs[g[-1]] if g else 0)
            tree_b += b"\0" * (24 + 8 + 16 * 32 - len(tree_b))
            taddr = w.write(tree_b)
            msgs.append(_msg(0x11, struct.pack("<QQ", taddr, haddr)))
        attr_blobs = []
        for k, v in node.attrs.items():
            head, mid, data = _attr_msg(w, k, v, gheap)
            if isinstance(data, tuple):
                body = head + mid + b"\0" * 16
                attr_blobs.append((len(msgs), len(head) + len(mid), data))
            else:
                body = head + mid + data
            if len(_pad8(body)) > 0xFFF0:
                raise H5Error(f"attribute {k!r} too large for a v1 object-header message")
            msgs.append(_msg(0x0C, body))
        payload = b"".join(msgs)
        hdr = struct.pack("<BBHII", 1, 0, len(msgs), 1, len(payload)) + b"\0" * 4
        at = w.write(hdr + payload)
        # remember where var-len descriptors live
        p = at + 16
        for i, m in enumerate(msgs):
            for (mi, inner, data) in attr_blobs:
                if mi == i:
                    vlen_fixups.append((p + 8 + inner, data[1], data[2]))
            p += len(m)
        return at

    # The root header must sit at 96 to mirror h5py; we instead point the superblock at wherever it lands.
    root_addr = emit(root)
    # global heap collection
    if gheap:
        body = bytearray()
        for i, blob in enumerate(gheap, 1):
            body += struct.pack("<HHIQ", i, 1, 0, len(blob)) + _pad8(blob)
        csize = max(4096, 16 + len(body) + 16)
        free = csize - 16 - len(body)
        body += struct.pack("<HHIQ", 0, 0, 0, free) + b"\0" * (free - 16)
        gaddr = w.write(b"GCOL" + struct.pack("<BBBBQ", 1, 0, 0, 0, csize) + bytes(body))
        for pos, ln, idx in vlen_fixups:
            w.patch(pos, struct.pack("<IQI", ln, gaddr, idx))
    w.align()
    eof = w.tell()
    sb = _SIG + struct.pack("<BBBBBBBBHHI", 0, 0, 0, 0, 0, 8, 8, 0, 4, 16, 0)
    sb += struct.pack("<QQQQ", 0, _UNDEF, eof, _UNDEF)
    sb += struct.pack("<QQII", 0, root_addr, 0, 0) + b"\0" * 16
    assert len(sb) == 96
    w.patch(0, sb)
    with open(path, "wb") as fh:
        fh.write(bytes(w.b))
