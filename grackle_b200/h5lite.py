"""Minimal pure-Python HDF5 reader/writer for the Grackle data-file dialect.

Test/bench infrastructure (there is no h5py or libhdf5 in this image).  Handles
exactly what Grackle's table files and yt golden-answer files use: superblock v0,
symbol-table groups (TREE/SNOD/local HEAP), version-1 object headers, contiguous
(or compact) un-filtered datasets, fixed-point / IEEE-float / fixed-length string
types of either byte order, and version-1 attribute messages.

The on-disk schema consumed by the reference is described in
initialize_cloudy_data.c:92-305 and initialize_UVbackground_data.c:50-330.
The writer emits the same dialect (big-endian f64/i64 like grackle_macros.h:58-62)
so the C++ loader in csrc/h5lite.cpp is exercised on both real and synthetic files.
"""
import struct

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF


class H5Object:
    def __init__(self, name):
        self.name = name
        self.attrs = {}
        self.children = {}   # groups
        self.data = None     # datasets (numpy array / bytes)

    def __getitem__(self, path):
        node = self
        for part in [p for p in path.split("/") if p]:
            node = node.children[part]
        return node

    def __contains__(self, path):
        try:
            self[path]
            return True
        except KeyError:
            return False

    def walk(self, prefix=""):
        for k, v in self.children.items():
            p = prefix + "/" + k
            yield p, v
            yield from v.walk(p)


# --------------------------------------------------------------------------
# reader
# --------------------------------------------------------------------------
class _Reader:
    def __init__(self, buf):
        self.b = buf
        if buf[:8] != b"\x89HDF\r\n\x1a\n":
            raise ValueError("not an HDF5 file")
        if buf[8] != 0:
            raise ValueError("only superblock version 0 is supported")
        if buf[13] != 8 or buf[14] != 8:
            raise ValueError("only 8-byte offsets/lengths are supported")
        self.base = struct.unpack_from("<Q", buf, 24)[0]

    def u(self, fmt, off):
        return struct.unpack_from("<" + fmt, self.b, off)

    # -- datatype message -> (numpy dtype or ('S',n)), size
    def parse_dtype(self, off):
        cv, b0, b1, b2, size = self.u("BBBBI", off)
        cls = cv & 0x0F
        if cls == 0:    # fixed point
            be = b0 & 1
            signed = (b0 >> 3) & 1
            return np.dtype((">" if be else "<") + ("i" if signed else "u") + str(size)), size
        if cls == 1:    # float
            be = b0 & 1
            return np.dtype((">" if be else "<") + "f" + str(size)), size
        if cls == 3:    # string
            return ("S", size), size
        if cls == 9:    # variable length (yt string attrs) - not decoded
            return ("VL", size), size
        raise ValueError("unsupported datatype class %d" % cls)

    def parse_dataspace(self, off):
        ver, rank, flags = self.u("BBB", off)
        if ver == 1:
            p = off + 8
        elif ver == 2:
            p = off + 4
        else:
            raise ValueError("dataspace version %d" % ver)
        dims = self.u("%dQ" % rank, p) if rank else ()
        return tuple(int(d) for d in dims)

    def decode(self, dt, dims, raw):
        n = 1
        for d in dims:
            n *= d
        if isinstance(dt, tuple):
            if dt[0] == "S":
                if n == 1 and not dims:
                    return raw[:dt[1]].split(b"\0")[0].decode("latin1")
                return [raw[i * dt[1]:(i + 1) * dt[1]].split(b"\0")[0].decode("latin1") for i in range(n)]
            return None
        arr = np.frombuffer(raw, dtype=dt, count=n).astype(dt.newbyteorder("="))
        return arr.reshape(dims) if dims else arr.reshape(())

    def messages(self, addr):
        """Yield (type, offset, size) for all header messages of a v1 object header."""
        ver, _, nmsg, _refc, hsize = self.u("BBHII", addr)
        if ver != 1:
            raise ValueError("object header version %d at %d" % (ver, addr))
        blocks = [(addr + 16, hsize)]
        count = 0
        while blocks and count < nmsg:
            p, left = blocks.pop(0)
            end = p + left
            while p + 8 <= end and count < nmsg:
                mtype, msize, _mflags = self.u("HHB", p)
                body = p + 8
                count += 1
                if mtype == 0x10:
                    coff, clen = self.u("QQ", body)
                    blocks.append((self.base + coff, clen))
                else:
                    yield mtype, body, msize
                p = body + msize

    def read_object(self, addr, name):
        obj = H5Object(name)
        dtype = dims = layout = None
        for mtype, body, msize in self.messages(addr):
            if mtype == 0x01:
                dims = self.parse_dataspace(body)
            elif mtype == 0x03:
                dtype, _ = self.parse_dtype(body)
            elif mtype == 0x08:
                layout = body
            elif mtype == 0x0C:
                self.read_attr(obj, body)
            elif mtype == 0x11:
                btree, heap = self.u("QQ", body)
                self.read_group(obj, self.base + btree, self.base + heap)
        if layout is not None and dtype is not None:
            obj.data = self.read_data(layout, dtype, dims or ())
            obj.dtype = dtype
        return obj

    def read_attr(self, obj, body):
        ver, _, nsz, dtsz, dssz = self.u("BBHHH", body)
        if ver != 1:
            return
        pad = lambda x: (x + 7) & ~7
        p = body + 8
        name = self.b[p:p + nsz].split(b"\0")[0].decode("latin1")
        p += pad(nsz)
        dt, _ = self.parse_dtype(p)
        p += pad(dtsz)
        dims = self.parse_dataspace(p)
        p += pad(dssz)
        n = 1
        for d in dims:
            n *= d
        size = (dt[1] if isinstance(dt, tuple) else dt.itemsize) * n
        obj.attrs[name] = self.decode(dt, dims, self.b[p:p + size])

    def read_data(self, layout, dt, dims):
        ver = self.b[layout]
        n = 1
        for d in dims:
            n *= d
        size = (dt[1] if isinstance(dt, tuple) else dt.itemsize) * n
        if ver == 3:
            cls = self.b[layout + 1]
            if cls == 1:      # contiguous
                addr, _sz = self.u("QQ", layout + 2)
                if addr == UNDEF:
                    return None
                return self.decode(dt, dims, self.b[self.base + addr:self.base + addr + size])
            if cls == 0:      # compact
                sz = self.u("H", layout + 2)[0]
                return self.decode(dt, dims, self.b[layout + 4:layout + 4 + sz])
            raise ValueError("chunked datasets are not supported")
        if ver in (1, 2):
            rank, cls = self.b[layout + 1], self.b[layout + 2]
            if cls != 1:
                raise ValueError("layout v%d class %d unsupported" % (ver, cls))
            addr = self.u("Q", layout + 8)[0]
            return self.decode(dt, dims, self.b[self.base + addr:self.base + addr + size])
        raise ValueError("layout version %d" % ver)

    def read_group(self, obj, btree, heap):
        assert self.b[heap:heap + 4] == b"HEAP"
        heap_data = self.base + self.u("Q", heap + 24)[0]
        self._walk_btree(obj, btree, heap_data)

    def _walk_btree(self, obj, node, heap_data):
        assert self.b[node:node + 4] == b"TREE", "bad B-tree node"
        _ntype, level, nent = self.u("BBH", node + 4)
        p = node + 24
        for i in range(nent):
            child = self.u("Q", p + 8 + i * 16)[0] + self.base
            if level > 0:
                self._walk_btree(obj, child, heap_data)
            else:
                self._read_snod(obj, child, heap_data)

    def _read_snod(self, obj, snod, heap_data):
        assert self.b[snod:snod + 4] == b"SNOD"
        nsym = self.u("H", snod + 6)[0]
        for i in range(nsym):
            e = snod + 8 + i * 40
            name_off, ohdr = self.u("QQ", e)
            s = heap_data + name_off
            name = self.b[s:self.b.index(b"\0", s)].decode("latin1")
            obj.children[name] = self.read_object(self.base + ohdr, name)


def read_file(path):
    with open(path, "rb") as f:
        buf = f.read()
    r = _Reader(buf)
    root_ohdr = struct.unpack_from("<Q", buf, 56 + 8)[0]
    return r.read_object(r.base + root_ohdr, "/")


# --------------------------------------------------------------------------
# writer (same dialect; used for the synthetic HM2012-schema table and for
# re-packing fixtures)
# --------------------------------------------------------------------------
class Dataset:
    def __init__(self, data, attrs=None):
        self.data = data          # numpy array (float64 / int64) or str
        self.attrs = attrs or {}


class Group(dict):
    """dict name -> Group | Dataset, plus .attrs"""
    def __init__(self, *a, attrs=None, **k):
        super().__init__(*a, **k)
        self.attrs = attrs or {}


def _pad8(b):
    return b + b"\0" * ((-len(b)) % 8)


def _dtype_msg(value):
    if isinstance(value, (str, bytes)):
        return struct.pack("<BBBBI", 0x13, 0, 0, 0, len(value) + 1)
    if np.asarray(value).dtype.kind in "iu":     # 64-bit signed big-endian
        return struct.pack("<BBBBIHH", 0x10, 0x09, 0, 0, 8, 0, 64)
    # IEEE f64 big-endian
    return struct.pack("<BBBBIHHBBBBI", 0x11, 0x21, 0x3F, 0, 8, 0, 64, 52, 11, 0, 52, 1023)


def _space_msg(value):
    dims = () if isinstance(value, (str, bytes)) else np.asarray(value).shape
    return struct.pack("<BBB5x", 1, len(dims), 0) + b"".join(struct.pack("<Q", d) for d in dims)


def _raw(value):
    if isinstance(value, str):
        return value.encode("latin1") + b"\0"
    if isinstance(value, bytes):
        return value + b"\0"
    a = np.asarray(value)
    return np.ascontiguousarray(a, dtype=">i8" if a.dtype.kind in "iu" else ">f8").tobytes()


def _msg(mtype, body):
    body = _pad8(body)
    return struct.pack("<HHB3x", mtype, len(body), 0) + body


def _attr_msg(name, value):
    nm = name.encode("latin1") + b"\0"
    dt, ds = _dtype_msg(value), _space_msg(value)
    body = (struct.pack("<BBHHH", 1, 0, len(nm), len(dt), len(ds)) + _pad8(nm) + _pad8(dt) +
            _pad8(ds) + _raw(value))
    return _msg(0x0C, body)


class _Writer:
    LEAF_K = 64     # up to 128 links per group in a single symbol-table node
    INT_K = 16

    def __init__(self):
        self.buf = bytearray(b"\0" * 96)

    def alloc(self, data):
        off = len(self.buf)
        self.buf += _pad8(bytes(data))
        return off

    def header(self, msgs):
        body = b"".join(msgs)
        return self.alloc(struct.pack("<BBHII4x", 1, 0, len(msgs), 1, len(body)) + body)

    def write_dataset(self, ds):
        raw = _raw(ds.data)
        addr = self.alloc(raw)
        msgs = [_msg(0x01, _space_msg(ds.data)), _msg(0x03, _dtype_msg(ds.data)),
                _msg(0x08, struct.pack("<BBQQ", 3, 1, addr, len(raw)))]
        msgs += [_attr_msg(k, v) for k, v in ds.attrs.items()]
        return self.header(msgs)

    def write_group(self, grp):
        """returns (object header address, btree address, heap address)"""
        names = sorted(grp.keys())
        assert len(names) <= 2 * self.LEAF_K, "too many entries for one symbol-table node"
        child = {}
        for n in names:
            node = grp[n]
            child[n] = self.write_dataset(node) if isinstance(node, Dataset) else self.write_group(node)[0]
        heap_data = bytearray(b"\0" * 8)         # offset 0 = empty string
        name_off = {}
        for n in names:
            name_off[n] = len(heap_data)
            heap_data += _pad8(n.encode("latin1") + b"\0")
        data_addr = self.alloc(heap_data)
        heap_addr = self.alloc(b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap_data), UNDEF, data_addr))
        snod = bytearray(b"SNOD" + struct.pack("<BBH", 1, 0, len(names)))
        for n in names:
            snod += struct.pack("<QQII16x", name_off[n], child[n], 0, 0)
        snod += b"\0" * (8 + 2 * self.LEAF_K * 40 - len(snod))
        snod_addr = self.alloc(snod)
        tree = bytearray(b"TREE" + struct.pack("<BBHQQ", 0, 0, 1, UNDEF, UNDEF))
        tree += struct.pack("<QQQ", 0, snod_addr, name_off[names[-1]] if names else 0)
        tree += b"\0" * (24 + (2 * self.INT_K + 1) * 8 + 2 * self.INT_K * 8 - len(tree))
        tree_addr = self.alloc(tree)
        msgs = [_msg(0x11, struct.pack("<QQ", tree_addr, heap_addr))]
        msgs += [_attr_msg(k, v) for k, v in getattr(grp, "attrs", {}).items()]
        return self.header(msgs), tree_addr, heap_addr


def write_file(path, root):
    """root: Group.  Writes a superblock-v0 HDF5 file."""
    w = _Writer()
    ohdr, tree_addr, heap_addr = w.write_group(root)
    sb = b"\x89HDF\r\n\x1a\n" + struct.pack("<BBBBBBBBHHI", 0, 0, 0, 0, 0, 8, 8, 0,
                                              w.LEAF_K, w.INT_K, 0)
    sb += struct.pack("<QQQQ", 0, UNDEF, len(w.buf), UNDEF)
    sb += struct.pack("<QQII", 0, ohdr, 1, 0) + struct.pack("<QQ", tree_addr, heap_addr)
    assert len(sb) == 96
    w.buf[:96] = sb
    with open(path, "wb") as f:
        f.write(bytes(w.buf))
