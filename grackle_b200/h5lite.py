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
