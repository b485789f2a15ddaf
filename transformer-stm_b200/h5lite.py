"""Dependency-free reader / writer for the subset of HDF5 that Keras-2 `model.save_weights('x.h5')` / `load_weights` use
(models/CvT(Par).py:489, models/CvT_test(Par).py:513).  h5py is not available in this environment.

What a Keras-2 weight file contains (keras/saving/hdf5_format.py, save_weights_to_hdf5_group):
  /                      attrs: layer_names [S], backend, keras_version  (long lists are split into layer_names0, layer_names1, ...)
  /<layer>/              attrs: weight_names [S]  (trainable weights first, then the non-trainable ones)
  /<layer>/<weight path> one contiguous float32 dataset per weight (the weight's name contains '/', so h5py nests groups)

Written here exactly the way libhdf5 writes it with default settings (what h5py.File(path, 'w') gives): superblock version 0,
version-1 object headers, old-style groups (symbol table message -> v1 B-tree -> SNOD nodes + local heap), contiguous layout,
fixed-length null-padded string attributes.  The reader additionally accepts superblock 2/3 files with version-2 object headers and
compact (link-message) groups, continuation blocks, version 1-3 attribute messages and version 1/2 dataspaces.  Chunked / compressed
datasets and dense (fractal-heap) groups are refused with a clear error: Keras does not produce them for weight files.

Format reference: "HDF5 File Format Specification Version 3.0" (The HDF Group), sections II-IV.
"""
from __future__ import annotations

import struct
from collections import OrderedDict
from typing import Dict, List, Optional, Tuple, Union

import numpy as np

SIGNATURE = b"\x89HDF\r\n\x1a\n"
UNDEF = 0xFFFFFFFFFFFFFFFF
LEAF_K, INTERNAL_K = 4, 16                      # libhdf5 defaults: symbol-table nodes hold 2*4 entries, B-tree nodes 2*16 children
HEADER_LIMIT = 64512                            # keras HDF5_OBJECT_HEADER_LIMIT: attributes larger than this are split


class H5Error(IOError):
    pass


class Group(OrderedDict):
    """name -> Group | np.ndarray; `.attrs` name -> np.ndarray / bytes / scalar."""

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self.attrs: "OrderedDict[str, object]" = OrderedDict()

    def require(self, path: str) -> "Group":
        g = self
        for part in [p for p in path.split("/") if p]:
            if part not in g:
                g[part] = Group()
            g = g[part]
            if not isinstance(g, Group):
                raise H5Error("%s is a dataset" % part)
        return g

    def get_path(self, path: str):
        g = self
        for part in [p for p in path.split("/") if p]:
            g = g[part]
        return g


def _pad8(n: int) -> int:
    return (n + 7) & ~7


# =====================================================================================================================
# writer
# =====================================================================================================================
def _dtype_msg(dt: np.dtype) -> bytes:
    dt = np.dtype(dt)
    if dt.kind == "f" and dt.itemsize in (4, 8):
        if dt.itemsize == 4:
            return struct.pack("<BBBBI", 0x11, 0x20, 31, 0, 4) + struct.pack("<HHBBBBI", 0, 32, 23, 8, 0, 23, 127)
        return struct.pack("<BBBBI", 0x11, 0x20, 63, 0, 8) + struct.pack("<HHBBBBI", 0, 64, 52, 11, 0, 52, 1023)
    if dt.kind in "iu":
        return struct.pack("<BBBBI", 0x10, 0x08 if dt.kind == "i" else 0x00, 0, 0, dt.itemsize) + struct.pack("<HH", 0, 8 * dt.itemsize)
    if dt.kind == "S":
        return struct.pack("<BBBBI", 0x13, 0x01, 0, 0, max(dt.itemsize, 1))          # null-padded ASCII, fixed length
    raise H5Error("unsupported dtype %s" % dt)


def _space_msg(shape: Tuple[int, ...]) -> bytes:
    return struct.pack("<BBB5x", 1, len(shape), 0) + b"".join(struct.pack("<Q", int(d)) for d in shape)


def _msg(mtype: int, body: bytes, flags: int = 0) -> bytes:
    body = body + b"\0" * (_pad8(len(body)) - len(body))
    return struct.pack("<HHB3x", mtype, len(body), flags) + body


def _attr_msg(name: str, value) -> bytes:
    a = np.asarray(value)
    if a.dtype.kind == "U":
        a = np.char.encode(a, "utf8")
    if a.dtype.kind == "O":
        raise H5Error("object arrays cannot be stored")
    if a.dtype.kind == "S" and a.dtype.itemsize == 0:
        a = a.astype("S1")
    nm = name.encode("utf8") + b"\0"
    dt, sp = _dtype_msg(a.dtype), _space_msg(a.shape)
    body = struct.pack("<BBHHH", 1, 0, len(nm), len(dt), len(sp))
    body += nm + b"\0" * (_pad8(len(nm)) - len(nm)) + dt + b"\0" * (_pad8(len(dt)) - len(dt)) + sp + b"\0" * (_pad8(len(sp)) - len(sp))
    body += a.tobytes()
    if len(body) > 65000:
        raise H5Error("attribute %s too large for an object-header message (%d bytes)" % (name, len(body)))
    return _msg(0x000C, body)


def _object_header(messages: List[bytes]) -> bytes:
    data = b"".join(messages)
    return struct.pack("<BBHII4x", 1, 0, len(messages), 1, len(data)) + data


class _Writer:
    def __init__(self):
        self.buf = bytearray(b"\0" * 96)            # superblock placeholder

    def alloc(self, data: bytes) -> int:
        pad = _pad8(len(self.buf)) - len(self.buf)
        self.buf += b"\0" * pad
        addr = len(self.buf)
        self.buf += data
        return addr

    def reserve(self, n: int) -> int:
        return self.alloc(b"\0" * n)

    def put(self, addr: int, data: bytes):
        self.buf[addr:addr + len(data)] = data

    def dataset(self, arr: np.ndarray) -> int:
        a = np.asarray(arr)                             # (ascontiguousarray would turn a scalar into shape (1,))
        if a.dtype.kind == "f" and a.dtype.itemsize not in (4, 8):
            a = a.astype(np.float32)
        raw = a.tobytes()
        daddr = self.alloc(raw) if raw else UNDEF
        msgs = [_msg(0x0001, _space_msg(a.shape)), _msg(0x0003, _dtype_msg(a.dtype), 1),
                _msg(0x0005, struct.pack("<BBBB", 2, 2, 2, 0)),
                _msg(0x0008, struct.pack("<BBQQ", 3, 1, daddr, len(raw)))]
        return self.alloc(_object_header(msgs))

    def group(self, g: Group) -> Tuple[int, int, int]:
        """Writes the group's children, then its heap / B-tree / symbol nodes / object header.  Returns (header, btree, heap) addresses."""
        children = []
        for name, child in g.items():
            if isinstance(child, Group):
                haddr, baddr, hpaddr = self.group(child)
                children.append((name, haddr, 1, baddr, hpaddr))
            else:
                children.append((name, self.dataset(np.asarray(child)), 0, 0, 0))
        children.sort(key=lambda c: c[0].encode("utf8"))
        # local heap: offset 0 = empty string (8 zero bytes), then the names, each padded to 8
        heap = bytearray(b"\0" * 8)
        offs = []
        for name, *_ in children:
            nm = name.encode("utf8") + b"\0"
            offs.append(len(heap))
            heap += nm + b"\0" * (_pad8(len(nm)) - len(nm))
        # libhdf5 wants room for a free block or the free-list head set to H5HL_FREE_NULL (1)
        heap_data = self.alloc(bytes(heap))
        heap_addr = self.alloc(b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap), 1, heap_data))
        # symbol table nodes of up to 2*LEAF_K entries; B-tree nodes of up to 2*INTERNAL_K children, as many levels as needed.
        # A node's key i is the largest name left of child i (heap offset; offset 0 = "" sorts before every name), key i+1 the largest in it.
        per = 2 * LEAF_K
        level = []                                                    # (address, heap offset of the largest name in the subtree)
        for i in range(0, len(children), per):
            idxs = list(range(i, min(i + per, len(children))))
            body = b"SNOD" + struct.pack("<BBH", 1, 0, len(idxs))
            for j in idxs:
                name, oaddr, ctype, baddr, hpaddr = children[j]
                body += struct.pack("<QQII", offs[j], oaddr, ctype, 0) + (struct.pack("<QQ", baddr, hpaddr) if ctype == 1 else b"\0" * 16)
            body += b"\0" * (8 + per * 40 - len(body))
            level.append((self.alloc(body), offs[idxs[-1]] if idxs else 0))
        depth = 0
        while True:
            fan = 2 * INTERNAL_K
            if not children:                                          # an empty group: a B-tree node without entries
                level = [(self.alloc(b"TREE" + struct.pack("<BBHQQ", 0, 0, 0, UNDEF, UNDEF) + b"\0" * ((fan + 1) * 8 + fan * 8)), 0)]
                break
            groups = [level[i:i + fan] for i in range(0, len(level), fan)]
            nxt, left_key = [], 0
            addrs = [self.reserve(24 + (fan + 1) * 8 + fan * 8) for _ in groups]
            for gi, grp in enumerate(groups):
                tree = b"TREE" + struct.pack("<BBHQQ", 0, depth, len(grp), addrs[gi - 1] if gi > 0 else UNDEF,
                                             addrs[gi + 1] if gi + 1 < len(groups) else UNDEF)
                tree += struct.pack("<Q", left_key)
                for caddr, ckey in grp:
                    tree += struct.pack("<QQ", caddr, ckey)
                self.put(addrs[gi], tree)
                left_key = grp[-1][1]
                nxt.append((addrs[gi], left_key))
            level, depth = nxt, depth + 1
            if len(level) == 1:
                break
        btree_addr = level[0][0]
        msgs = [_msg(0x0011, struct.pack("<QQ", btree_addr, heap_addr))]
        for k, v in g.attrs.items():
            msgs.append(_attr_msg(k, v))
        return self.alloc(_object_header(msgs)), btree_addr, heap_addr


def write_h5(path: str, root: Group):
    w = _Writer()
    haddr, baddr, hpaddr = w.group(root)
    eof = _pad8(len(w.buf))
    w.buf += b"\0" * (eof - len(w.buf))
    sb = SIGNATURE + struct.pack("<BBBBBBBBHHI", 0, 0, 0, 0, 0, 8, 8, 0, LEAF_K, INTERNAL_K, 0)
    sb += struct.pack("<QQQQ", 0, UNDEF, eof, UNDEF)
    sb += struct.pack("<QQII", 0, haddr, 1, 0) + struct.pack("<QQ", baddr, hpaddr)
    assert len(sb) == 96
    w.put(0, sb)
    with open(path, "wb") as f:
        f.write(bytes(w.buf))


# =====================================================================================================================
# reader
# =====================================================================================================================
class _Reader:
    def __init__(self, data: bytes):
        self.d = data
        self.base = 0
        if data[:8] != SIGNATURE:
            # the superblock may sit at 512, 1024, 2048, ... (user block); Keras never writes one, but be correct
            off = 512
            while off < len(data) and data[off:off + 8] != SIGNATURE:
                off *= 2
            if off >= len(data):
                raise H5Error("not an HDF5 file")
            self.base = off
        ver = data[self.base + 8]
        if ver in (0, 1):
            o = self.base + 8
            so, sl = data[o + 5], data[o + 6]
            if (so, sl) != (8, 8):
                raise H5Error("only 8-byte offsets / lengths are supported (got %d / %d)" % (so, sl))
            o += 16 + (4 if ver == 1 else 0)
            base_addr = struct.unpack_from("<Q", data, o)[0]
            self.base += 0 if base_addr == 0 else 0
            o += 32                                                  # base, free-space, eof, driver
            self.root = struct.unpack_from("<Q", data, o + 8)[0]
        elif ver in (2, 3):
            o = self.base + 9
            so, sl = data[o], data[o + 1]
            if (so, sl) != (8, 8):
                raise H5Error("only 8-byte offsets / lengths are supported")
            self.root = struct.unpack_from("<Q", data, o + 3 + 24)[0]
        else:
            raise H5Error("unsupported superblock version %d" % ver)

    # ---- object headers -> list of (type, flags, body) ----
    def messages(self, addr: int) -> List[Tuple[int, int, bytes]]:
        d, a = self.d, self.base + addr
        out: List[Tuple[int, int, bytes]] = []
        if d[a:a + 4] == b"OHDR":
            if d[a + 4] != 2:
                raise H5Error("object header version %d" % d[a + 4])
            flags = d[a + 5]
            o = a + 6
            if flags & 0x20:
                o += 16
            if flags & 0x10:
                o += 4
            szb = 1 << (flags & 3)
            size = int.from_bytes(d[o:o + szb], "little"); o += szb
            blocks = [(o, size)]
            while blocks:
                o, size = blocks.pop(0)
                end = o + size
                while o + 4 <= end:
                    mt, ms, mf = d[o], struct.unpack_from("<H", d, o + 1)[0], d[o + 3]
                    o += 4 + (2 if flags & 4 else 0)
                    body = d[o:o + ms]
                    o += ms
                    if mt == 0x10:
                        ca, cl = struct.unpack_from("<QQ", body)
                        if d[self.base + ca:self.base + ca + 4] != b"OCHK":
                            raise H5Error("bad continuation block")
                        blocks.append((self.base + ca + 4, cl - 8))      # minus signature and checksum
                    elif mt != 0:
                        out.append((mt, mf, body))
            return out
        ver, _, nmsg, _, hsize = struct.unpack_from("<BBHII", d, a)
        if ver != 1:
            raise H5Error("object header version %d at %d" % (ver, addr))
        blocks = [(a + 16, hsize)]
        while blocks and len(out) < 100000:
            o, size = blocks.pop(0)
            end = o + size
            while o + 8 <= end:
                mt, ms, mf = struct.unpack_from("<HHB", d, o)
                body = d[o + 8:o + 8 + ms]
                o += 8 + ms
                if mt == 0x10:
                    ca, cl = struct.unpack_from("<QQ", body)
                    blocks.append((self.base + ca, cl))
                elif mt != 0:
                    out.append((mt, mf, body))
        return out

    # ---- datatype / dataspace ----
    @staticmethod
    def parse_dtype(b: bytes) -> Tuple[Optional[np.dtype], int]:
        cls, ver = b[0] & 0x0F, b[0] >> 4
        size = struct.unpack_from("<I", b, 4)[0]
        if cls == 1:
            if b[1] & 1:
                return np.dtype(">f%d" % size), size
            return np.dtype("<f%d" % size), size
        if cls == 0:
            signed = bool(b[1] & 0x08)
            return np.dtype(("<" if not (b[1] & 1) else ">") + ("i" if signed else "u") + str(size)), size
        if cls == 3:
            return np.dtype("S%d" % size), size
        return None, size                      # variable-length strings, compounds, ...: not needed for weight files

    @staticmethod
    def parse_space(b: bytes) -> Tuple[int, ...]:
        ver, rank, flags = b[0], b[1], b[2]
        if ver == 1:
            o = 8
        elif ver == 2:
            o = 4
            if b[3] == 2:
                return (0,)
        else:
            raise H5Error("dataspace version %d" % ver)
        return tuple(struct.unpack_from("<Q", b, o + 8 * i)[0] for i in range(rank))

    def parse_attr(self, b: bytes):
        ver = b[0]
        nsz, dsz, ssz = struct.unpack_from("<HHH", b, 2)
        if ver == 1:
            o = 8
            name = b[o:o + nsz].split(b"\0")[0].decode("utf8"); o += _pad8(nsz)
            dt = b[o:o + dsz]; o += _pad8(dsz)
            sp = b[o:o + ssz]; o += _pad8(ssz)
        elif ver in (2, 3):
            o = 8 + (1 if ver == 3 else 0)
            name = b[o:o + nsz].split(b"\0")[0].decode("utf8"); o += nsz
            dt = b[o:o + dsz]; o += dsz
            sp = b[o:o + ssz]; o += ssz
        else:
            raise H5Error("attribute version %d" % ver)
        dtype, size = self.parse_dtype(dt)
        shape = self.parse_space(sp)
        if dtype is None:
            return name, None
        n = int(np.prod(shape, dtype=np.int64)) if shape else 1
        val = np.frombuffer(b, dtype=dtype, count=n, offset=o).reshape(shape).copy()
        if dtype.kind == "S":
            val = np.char.rstrip(val, b"\0") if val.shape else val
        return name, (val if val.shape else val[()])

    # ---- groups ----
    def heap_string(self, heap_addr: int, off: int) -> str:
        d, a = self.d, self.base + heap_addr
        if d[a:a + 4] != b"HEAP":
            raise H5Error("bad local heap")
        data_addr = struct.unpack_from("<Q", d, a + 24)[0]
        s = self.base + data_addr + off
        return d[s:d.index(b"\0", s)].decode("utf8")

    def btree_links(self, btree: int, heap: int, out: List[Tuple[str, int]]):
        d, a = self.d, self.base + btree
        if d[a:a + 4] == b"SNOD":
            n = struct.unpack_from("<H", d, a + 6)[0]
            for i in range(n):
                noff, oaddr = struct.unpack_from("<QQ", d, a + 8 + 40 * i)
                out.append((self.heap_string(heap, noff), oaddr))
            return
        if d[a:a + 4] != b"TREE":
            raise H5Error("bad group B-tree node")
        ntype, level, used = struct.unpack_from("<BBH", d, a + 4)
        if ntype != 0:
            raise H5Error("chunked datasets are not supported")
        o = a + 24
        for i in range(used):
            child = struct.unpack_from("<Q", d, o + 8)[0]
            self.btree_links(child, heap, out)
            o += 16

    def read_object(self, addr: int, depth: int = 0):
        if depth > 64:
            raise H5Error("group nesting too deep")
        msgs = self.messages(addr)
        types = {t for t, _, _ in msgs}
        attrs = OrderedDict()
        for t, _, b in msgs:
            if t == 0x0C:
                k, v = self.parse_attr(b)
                if v is not None:
                    attrs[k] = v
        if 0x11 in types or 0x06 in types or 0x02 in types:
            g = Group()
            g.attrs = attrs
            links: List[Tuple[str, int]] = []
            for t, _, b in msgs:
                if t == 0x11:
                    bt, hp = struct.unpack_from("<QQ", b)
                    self.btree_links(bt, hp, links)
                elif t == 0x06:
                    lf = b[1]
                    o = 2
                    ltype = 0
                    if lf & 0x08:
                        ltype = b[o]; o += 1
                    if lf & 0x04:
                        o += 8
                    if lf & 0x10:
                        o += 1
                    lsz = 1 << (lf & 3)
                    ln = int.from_bytes(b[o:o + lsz], "little"); o += lsz
                    name = b[o:o + ln].decode("utf8"); o += ln
                    if ltype == 0:
                        links.append((name, struct.unpack_from("<Q", b, o)[0]))
                elif t == 0x02:
                    fh = struct.unpack_from("<Q", b, 2 + (8 if b[1] & 1 else 0))[0]
                    if fh != UNDEF:
                        raise H5Error("dense (fractal heap) groups are not supported; re-save the file with default h5py settings")
            for name, oaddr in links:
                g[name] = self.read_object(oaddr, depth + 1)
            return g
        # dataset
        shape = dtype = None
        data = None
        for t, _, b in msgs:
            if t == 0x01:
                shape = self.parse_space(b)
            elif t == 0x03:
                dtype, _ = self.parse_dtype(b)
            elif t == 0x08:
                ver, cls = b[0], b[1]
                if ver != 3 and ver != 4:
                    raise H5Error("data layout version %d" % ver)
                if cls == 1:
                    daddr, dsize = struct.unpack_from("<QQ", b, 2)
                    data = ("contig", daddr, dsize)
                elif cls == 0:
                    sz = struct.unpack_from("<H", b, 2)[0]
                    data = ("compact", b[4:4 + sz])
                else:
                    raise H5Error("chunked / virtual dataset layouts are not supported (Keras writes contiguous weights)")
        if shape is None or dtype is None or data is None:
            raise H5Error("object at %d is neither a group nor a readable dataset" % addr)
        n = int(np.prod(shape, dtype=np.int64)) if shape else 1
        if data[0] == "contig":
            if data[1] == UNDEF or n == 0:
                arr = np.zeros(shape, dtype)
            else:
                arr = np.frombuffer(self.d, dtype=dtype, count=n, offset=self.base + data[1]).reshape(shape).copy()
        else:
            arr = np.frombuffer(data[1], dtype=dtype, count=n).reshape(shape).copy()
        return arr


def read_h5(path: str) -> Group:
    with open(path, "rb") as f:
        data = f.read()
    r = _Reader(data)
    root = r.read_object(r.root)
    if not isinstance(root, Group):
        raise H5Error("root object is not a group")
    return root


# =====================================================================================================================
# Keras-2 weight-file layer on top
# =====================================================================================================================
def _split_attr(g: Group, name: str, values: List[bytes]):
    """keras save_attributes_to_hdf5_group: one attribute, or name0, name1, ... when it would exceed the object-header limit."""
    arr = np.asarray(values, dtype="S%d" % max([len(v) for v in values] + [1]))
    if arr.nbytes <= HEADER_LIMIT:
        g.attrs[name] = arr
        return
    parts = 1
    chunks = np.array_split(arr, parts)
    while any(c.nbytes > HEADER_LIMIT for c in chunks):
        parts += 1
        chunks = np.array_split(arr, parts)
    for i, c in enumerate(chunks):
        g.attrs["%s%d" % (name, i)] = c


def _joined_attr(g: Group, name: str) -> List[str]:
    """keras load_attributes_from_hdf5_group."""
    if name in g.attrs:
        vals = list(np.atleast_1d(g.attrs[name]))
    else:
        vals, i = [], 0
        while "%s%d" % (name, i) in g.attrs:
            vals.extend(np.atleast_1d(g.attrs["%s%d" % (name, i)]))
            i += 1
    return [v.decode("utf8") if isinstance(v, (bytes, np.bytes_)) else str(v) for v in vals]


def save_keras_weights(path: str, layers: List[Tuple[str, List[Tuple[str, np.ndarray]]]], keras_version: str = "2.11.0"):
    """layers: [(layer_name, [(weight_name, array), ...])] in model.layers order; layers without weights are listed with []."""
    root = Group()
    _split_attr(root, "layer_names", [n.encode("utf8") for n, _ in layers])
    root.attrs["backend"] = np.bytes_(b"tensorflow")
    root.attrs["keras_version"] = np.bytes_(keras_version.encode("utf8"))
    for lname, weights in layers:
        g = root.require(lname)
        if weights:
            _split_attr(g, "weight_names", [wn.encode("utf8") for wn, _ in weights])
        else:
            g.attrs["weight_names"] = np.zeros((0,), "S1")
        for wn, arr in weights:
            parts = wn.split("/")
            sub = g.require("/".join(parts[:-1])) if len(parts) > 1 else g
            sub[parts[-1]] = np.asarray(arr, dtype=np.float32)
    write_h5(path, root)


def load_keras_weights(path: str) -> "OrderedDict[str, List[Tuple[str, np.ndarray]]]":
    """-> {layer_name: [(weight_name, array), ...]} in the file's order (trainable first, non-trainable last, as Keras wrote them)."""
    root = read_h5(path)
    if "model_weights" in root and "layer_names" not in root.attrs and not any(k.startswith("layer_names") for k in root.attrs):
        root = root["model_weights"]            # a full model.save('x.h5') file keeps the same layout one level down
    out: "OrderedDict[str, List[Tuple[str, np.ndarray]]]" = OrderedDict()
    for lname in _joined_attr(root, "layer_names"):
        g = root[lname]
        out[lname] = [(wn, np.asarray(g.get_path(wn))) for wn in _joined_attr(g, "weight_names")]
    return out
