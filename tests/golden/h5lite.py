"""Minimal pure-Python reader for the NetCDF-4 / HDF5 files the reference ships (test infrastructure, used only by
make_mesh_fixtures.py in the container that has /root/reference; nothing on the GPU box imports it).

Covers exactly what those files use (HDF5 File Format Specification 3.0):
  superblock v2/v3 -> object headers v2 ('OHDR', continuation chunks 'OCHK') -> messages
  link info (0x02) with dense storage: fractal heap ('FRHP', root direct block 'FHDB' or one-level indirect block 'FHIB') holding link messages,
  compact link messages (0x06), dataspace (0x01) v1/v2, datatype (0x03) fixed-point / floating-point / string,
  data layout (0x08) v3/v4: compact, contiguous, chunked through a v1 B-tree ('TREE', node type 1) WITHOUT filters,
  attributes (0x0C) v1-v3 of numeric / fixed-string type.
Anything else raises H5Unsupported naming the structure, so that a new file that needs more says exactly what.
"""
import struct

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF


class H5Unsupported(Exception):
    pass


class _Buf:
    def __init__(self, d, p=0):
        self.d, self.p = d, p

    def u(self, n):
        v = int.from_bytes(self.d[self.p:self.p + n], "little")
        self.p += n
        return v

    def raw(self, n):
        v = self.d[self.p:self.p + n]
        self.p += n
        return v

    def skip(self, n):
        self.p += n


class Dataset:
    def __init__(self, f, name, shape, dtype, layout, attrs):
        self.f, self.name, self.shape, self.dtype, self.layout, self.attrs = f, name, shape, dtype, layout, attrs

    def read(self):
        return self.f._read_data(self)


class File:
    def __init__(self, path):
        with open(path, "rb") as fh:
            self.d = fh.read()
        d = self.d
        if d[:8] != b"\x89HDF\r\n\x1a\n":
            raise H5Unsupported("not an HDF5 file (signature at offset 0)")
        ver = d[8]
        if ver not in (2, 3):
            raise H5Unsupported(f"superblock version {ver}")
        self.O, self.L = d[9], d[10]
        if (self.O, self.L) != (8, 8):
            raise H5Unsupported(f"offset/length sizes {self.O}/{self.L}")
        b = _Buf(d, 12)
        self.base = b.u(8); b.u(8); b.u(8)
        self.root = b.u(8)
        self.datasets = {}
        self.root_attrs = {}
        self._walk_group(self.root, "")

    # ---- object headers -------------------------------------------------------------------------------------------
    def _messages(self, addr):
        d = self.d
        if d[addr:addr + 4] != b"OHDR":
            raise H5Unsupported(f"object header v1 (no OHDR signature at {addr})")
        b = _Buf(d, addr + 4)
        if b.u(1) != 2:
            raise H5Unsupported("object header version")
        flags = b.u(1)
        if flags & 0x20:
            b.skip(16)
        if flags & 0x10:
            b.skip(4)
        size0 = b.u(1 << (flags & 3))
        track_order = bool(flags & 0x04)
        out = []
        chunks = [(b.p, b.p + size0)]
        while chunks:
            p, end = chunks.pop(0)
            b = _Buf(d, p)
            while b.p + 4 <= end:
                typ = b.u(1); size = b.u(2); mflags = b.u(1)
                if track_order:
                    b.skip(2)
                body = b.raw(size)
                if typ == 0x10:   # continuation
                    cb = _Buf(body)
                    caddr, clen = cb.u(8), cb.u(8)
                    if d[caddr:caddr + 4] != b"OCHK":
                        raise H5Unsupported("continuation block without OCHK")
                    chunks.append((caddr + 4, caddr + clen - 4))
                elif typ != 0:
                    out.append((typ, mflags, body))
        return out

    # ---- groups ---------------------------------------------------------------------------------------------------
    def _parse_link(self, b):
        ver = b.u(1)
        if ver != 1:
            return None
        fl = b.u(1)
        ltype = b.u(1) if fl & 0x08 else 0
        if fl & 0x04:
            b.skip(8)
        if fl & 0x10:
            b.skip(1)
        ln = b.u(1 << (fl & 3))
        name = b.raw(ln).decode("utf-8")
        if ltype != 0:
            raise H5Unsupported(f"link type {ltype} ({name})")
        return name, b.u(8)

    def _heap_links(self, heap_addr):
        d = self.d
        if d[heap_addr:heap_addr + 4] != b"FRHP":
            raise H5Unsupported("fractal heap header")
        b = _Buf(d, heap_addr + 5)
        b.u(2); filt_len = b.u(2); hflags = b.u(1)
        if filt_len:
            raise H5Unsupported("filtered fractal heap")
        b.u(4)
        b.skip(8 * 12)   # huge id, huge btree, free space, fs manager, managed space, allocated, iterator offset, n managed, huge size, n huge, tiny size, n tiny
        width = b.u(2); start = b.u(8); max_direct = b.u(8); max_heap_bits = b.u(2); b.u(2)
        root = b.u(8); cur_rows = b.u(2)
        off_bytes = (max_heap_bits + 7) // 8
        links = []

        def direct(addr, size):
            if d[addr:addr + 4] != b"FHDB":
                raise H5Unsupported("fractal heap direct block")
            p = addr + 5 + 8 + off_bytes + (4 if hflags & 0x02 else 0)
            bb = _Buf(d, p)
            while bb.p < addr + size and d[bb.p] == 1:
                r = self._parse_link(bb)
                if r is None:
                    break
                links.append(r)

        if root == UNDEF:
            return links
        if cur_rows == 0:
            direct(root, start)
        else:
            if d[root:root + 4] != b"FHIB":
                raise H5Unsupported("fractal heap indirect block")
            max_direct_rows = (max_direct.bit_length() - 1) - (start.bit_length() - 1) + 2
            if cur_rows > max_direct_rows:
                raise H5Unsupported("fractal heap with nested indirect blocks")
            bb = _Buf(d, root + 5 + 8 + off_bytes)
            for row in range(cur_rows):
                size = start if row < 2 else start << (row - 1)
                for _ in range(width):
                    a = bb.u(8)
                    if a != UNDEF:
                        direct(a, size)
        return links

    def _walk_group(self, addr, prefix):
        msgs = self._messages(addr)
        links = []
        is_dataset = any(t == 0x08 for t, _, _ in msgs)
        if is_dataset:
            return self._make_dataset(prefix, msgs)
        for typ, _, body in msgs:
            if typ == 0x06:
                links.append(self._parse_link(_Buf(body)))
            elif typ == 0x02:
                b = _Buf(body)
                b.u(1); fl = b.u(1)
                if fl & 1:
                    b.skip(8)
                heap = b.u(8)
                if heap != UNDEF:
                    links += self._heap_links(heap)
            elif typ == 0x0C and prefix == "":
                k, v = self._attribute(body)
                self.root_attrs[k] = v
        for name, a in links:
            self._walk_group(a, (prefix + "/" if prefix else "") + name)

    # ---- datasets -------------------------------------------------------------------------------------------------
    @staticmethod
    def _dataspace(body):
        b = _Buf(body)
        ver = b.u(1); rank = b.u(1); fl = b.u(1)
        if ver == 1:
            b.skip(5)
        elif ver == 2:
            b.skip(1)
        else:
            raise H5Unsupported(f"dataspace version {ver}")
        return tuple(b.u(8) for _ in range(rank))

    @staticmethod
    def _datatype(body):
        b = _Buf(body)
        cv = b.u(1); bits = b.raw(3); size = b.u(4)
        cls = cv & 0x0F
        if cls == 0:
            signed = bool(bits[0] & 0x08)
            return np.dtype(("<" if not bits[0] & 1 else ">") + ("i" if signed else "u") + str(size))
        if cls == 1:
            return np.dtype(("<" if not bits[0] & 1 else ">") + "f" + str(size))
        if cls == 3:
            return np.dtype("S" + str(size))
        return None   # variable-length / reference / compound ... (dimension lists): skipped by the caller

    def _attribute(self, body):
        b = _Buf(body)
        ver = b.u(1)
        fl = b.u(1)
        nlen, tlen, slen = b.u(2), b.u(2), b.u(2)
        if ver == 3:
            b.skip(1)
        pad = (lambda n: (n + 7) & ~7) if ver == 1 else (lambda n: n)
        name = b.raw(pad(nlen))[:nlen].rstrip(b"\0").decode()
        if fl & 3:
            return name, None
        dt = self._datatype(b.raw(pad(tlen))[:tlen])
        shape = self._dataspace(b.raw(pad(slen))[:slen]) if slen else ()
        if dt is None:
            return name, None
        n = int(np.prod(shape)) if shape else 1
        v = np.frombuffer(b.raw(n * dt.itemsize), dtype=dt)
        if dt.kind == "S":
            return name, v[0].rstrip(b"\0").decode("utf-8", "replace") if n == 1 else [x.decode() for x in v]
        return name, (v[0] if shape == () else v.copy())

    def _make_dataset(self, name, msgs):
        shape = dtype = layout = None
        attrs = {}
        for typ, _, body in msgs:
            if typ == 0x01:
                shape = self._dataspace(body)
            elif typ == 0x03:
                dtype = self._datatype(body)
            elif typ == 0x0B:
                raise H5Unsupported(f"filter pipeline on dataset {name}")
            elif typ == 0x08:
                b = _Buf(body)
                ver = b.u(1)
                if ver not in (3, 4):
                    raise H5Unsupported(f"layout message version {ver}")
                cls = b.u(1)
                if cls == 0:
                    n = b.u(2)
                    layout = ("compact", b.raw(n))
                elif cls == 1:
                    layout = ("contiguous", b.u(8), b.u(8))
                elif cls == 2 and ver == 3:
                    rank = b.u(1)
                    bt = b.u(8)
                    dims = [b.u(4) for _ in range(rank)]
                    layout = ("chunked", bt, dims)
                else:
                    raise H5Unsupported(f"layout class {cls} version {ver} on {name}")
            elif typ == 0x0C:
                try:
                    k, v = self._attribute(body)
                    attrs[k] = v
                except Exception:
                    pass
        if dtype is None:
            return
        self.datasets[name] = Dataset(self, name, shape, dtype, layout, attrs)

    def _chunks(self, addr, rank1):
        """leaves of a v1 B-tree of raw-data chunks: (offsets, size, filter mask, address)"""
        d = self.d
        if d[addr:addr + 4] != b"TREE":
            raise H5Unsupported("v1 B-tree node")
        b = _Buf(d, addr + 4)
        ntype, level, n = b.u(1), b.u(1), b.u(2)
        if ntype != 1:
            raise H5Unsupported("B-tree node type")
        b.skip(16)
        out = []
        for _ in range(n):
            size = b.u(4); mask = b.u(4)
            offs = [b.u(8) for _ in range(rank1)]
            child = b.u(8)
            if level == 0:
                out.append((offs, size, mask, child))
            else:
                out += self._chunks(child, rank1)
        return out

    def _read_data(self, ds):
        n = int(np.prod(ds.shape)) if ds.shape else 1
        lay = ds.layout
        if lay[0] == "compact":
            return np.frombuffer(lay[1][:n * ds.dtype.itemsize], dtype=ds.dtype).reshape(ds.shape).copy()
        if lay[0] == "contiguous":
            if lay[1] == UNDEF:
                return np.zeros(ds.shape, ds.dtype)
            a = self.base + lay[1]
            return np.frombuffer(self.d[a:a + n * ds.dtype.itemsize], dtype=ds.dtype).reshape(ds.shape).copy()
        _, bt, dims = lay
        cdims = dims[:-1]
        out = np.zeros(ds.shape, ds.dtype)
        if bt == UNDEF:
            return out
        for offs, size, mask, a in self._chunks(bt, len(dims)):
            if mask:
                raise H5Unsupported("filtered chunk")
            c = np.frombuffer(self.d[a:a + size], dtype=ds.dtype)[:int(np.prod(cdims))].reshape(cdims)
            sl = tuple(slice(o, min(o + c_, s)) for o, c_, s in zip(offs, cdims, ds.shape))
            out[sl] = c[tuple(slice(0, s.stop - s.start) for s in sl)]
        return out

    def __getitem__(self, name):
        return self.datasets[name].read()

    def keys(self):
        return list(self.datasets)
