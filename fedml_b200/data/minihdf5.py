"""A small HDF5 reader and writer in pure Python + numpy, with the slice of the h5py interface the dataset loaders use.

Why: the federated benchmarks of the reference ship as TensorFlow-Federated ``.h5`` exports - ``fed_emnist_*.h5``,
``fed_cifar100_*.h5``, ``shakespeare_*.h5``, ``stackoverflow_*.h5`` (reference data/FederatedEMNIST/data_loader.py:30-75,
data/fed_cifar100/data_loader.py:33-70, data/fed_shakespeare/data_loader.py:30-70, data/stackoverflow_nwp/dataset.py:30-56) -
and the reference opens them with h5py, which is a C extension that is not part of this image.  Those files only use a
narrow part of the format, which is what is implemented here directly on the bytes (through ``mmap``):

* superblock versions 0-3, with a user block in front (the signature is searched at 0, 512, 1024, ...);
* object headers version 1 and 2, with continuation blocks;
* groups stored as symbol tables (B-tree v1 + symbol nodes + local heap - what h5py writes by default) and as compact link
  messages; "dense" groups (fractal heap, only written with ``libver='latest'``) raise ``Hdf5Unsupported``;
* datasets: compact, contiguous and chunked (B-tree v1 chunk index, or the version-4 "single chunk" / "implicit" index) with the
  deflate, shuffle and fletcher32 filters;
* element types: integers and floats of either byte order, fixed-length strings, variable-length strings / sequences
  (global heap), enums (read as their base integer, which covers h5py booleans) and compound records;
* attributes (versions 1-3) of those types.

``File`` / ``Group`` / ``Dataset`` follow h5py: ``f["examples"]["client"]["pixels"][()]``, ``.keys()``, ``in``, ``.shape``,
``.dtype``, ``.attrs``, ``with File(path) as f``.  Variable-length strings come back as ``bytes`` objects like in h5py (callers
``.decode("utf8")``).

``write_hdf5(path, tree)`` is the inverse for nested dicts of arrays (symbol-table groups with multi-level B-trees, contiguous
or chunked+deflate data sets, variable-length strings): it is what ``tools/export_tff_h5.py`` and the tests use to produce files
in the TFF layout.
"""
from __future__ import annotations

import mmap
import struct
import zlib
from typing import Dict, Iterator, List, Optional, Tuple

import numpy as np

SIGNATURE = b"\x89HDF\r\n\x1a\n"
UNDEF = 0xFFFFFFFFFFFFFFFF


class Hdf5Error(ValueError):
    pass


class Hdf5Unsupported(NotImplementedError):
    pass


# message types
M_DATASPACE, M_LINK_INFO, M_DATATYPE, M_FILL_OLD, M_FILL, M_LINK, M_LAYOUT, M_FILTERS, M_ATTRIBUTE, M_CONT, M_SYMTAB = \
    0x1, 0x2, 0x3, 0x4, 0x5, 0x6, 0x8, 0xB, 0xC, 0x10, 0x11


class _Type:
    """Decoded datatype message: numpy dtype for plain types, or a marker for variable-length data."""

    def __init__(self, dtype: Optional[np.dtype], size: int, vlen: Optional[str] = None, base: Optional["_Type"] = None):
        self.dtype, self.size, self.vlen, self.base = dtype, size, vlen, base


class _Image:
    """The file bytes plus the parsing routines that need the superblock's offset / length sizes."""

    def __init__(self, path: str):
        self.path = path
        self.fh = open(path, "rb")
        try:
            self.buf = mmap.mmap(self.fh.fileno(), 0, access=mmap.ACCESS_READ)
        except ValueError as exc:
            self.fh.close()
            raise Hdf5Error(f"{path}: empty file") from exc
        self._gheaps: Dict[int, Dict[int, Tuple[int, int]]] = {}
        self._superblock()

    def close(self):
        if self.buf is not None:
            self.buf.close()
            self.fh.close()
            self.buf = None

    # ---- primitives
    def u(self, off: int, n: int) -> int:
        return int.from_bytes(self.buf[off:off + n], "little")

    def addr(self, off: int) -> int:
        a = self.u(off, self.O)
        return UNDEF if a == (1 << (8 * self.O)) - 1 else a + self.base

    def length(self, off: int) -> int:
        return self.u(off, self.L)

    def _superblock(self):
        n, pos = len(self.buf), 0
        while pos + 8 <= n and self.buf[pos:pos + 8] != SIGNATURE:
            pos = 512 if pos == 0 else pos * 2
        if pos + 8 > n:
            raise Hdf5Error(f"{self.path}: no HDF5 signature")
        ver = self.buf[pos + 8]
        self.base = 0
        if ver in (0, 1):
            self.O, self.L = self.buf[pos + 13], self.buf[pos + 14]
            self.leaf_k, self.internal_k = self.u(pos + 16, 2), self.u(pos + 18, 2)
            p = pos + 24 + (4 if ver == 1 else 0)
            base = self.u(p, self.O)
            self.base = base if base != (1 << (8 * self.O)) - 1 else pos
            p += 4 * self.O                                    # base, free-space, end-of-file, driver-info addresses
            self.root_header = self.addr(p + self.O)           # symbol table entry: name offset, header address, ...
        elif ver in (2, 3):
            self.O, self.L = self.buf[pos + 9], self.buf[pos + 10]
            self.base = self.u(pos + 12, self.O)
            self.root_header = self.addr(pos + 12 + 3 * self.O)
        else:
            raise Hdf5Unsupported(f"{self.path}: superblock version {ver}")
        if self.O not in (4, 8) or self.L not in (4, 8):
            raise Hdf5Error(f"{self.path}: offset/length sizes {self.O}/{self.L}")

    # ---- object headers
    def messages(self, header: int) -> List[Tuple[int, int, int]]:
        """(type, offset of the body, size of the body) of every message of the object header at ``header``."""
        if self.buf[header:header + 4] == b"OHDR":
            return self._messages_v2(header)
        if self.buf[header] != 1:
            raise Hdf5Error(f"{self.path}: no object header at {header:#x}")
        total = self.u(header + 2, 2)
        blocks = [(header + 16, self.u(header + 8, 4))]
        out = []
        while blocks and len(out) < total:
            p, size = blocks.pop(0)
            end = p + size
            while p + 8 <= end and len(out) < total:
                mtype, msize = self.u(p, 2), self.u(p + 2, 2)
                body = p + 8
                if mtype == M_CONT:
                    blocks.append((self.addr(body), self.length(body + self.O)))
                out.append((mtype, body, msize))
                p = body + msize
        return out

    def _messages_v2(self, header: int) -> List[Tuple[int, int, int]]:
        flags = self.buf[header + 5]
        p = header + 6
        if flags & 0x20:
            p += 16
        if flags & 0x10:
            p += 4
        nsz = 1 << (flags & 3)
        size = self.u(p, nsz)
        p += nsz
        blocks = [(p, size)]
        out = []
        tracked = 2 if flags & 0x4 else 0
        while blocks:
            p, size = blocks.pop(0)
            end = p + size
            while p + 4 + tracked <= end:
                mtype, msize = self.buf[p], self.u(p + 1, 2)
                body = p + 4 + tracked
                if body + msize > end:
                    break
                if mtype == M_CONT:
                    a, ln = self.addr(body), self.length(body + self.O)
                    blocks.append((a + 4, ln - 8))              # "OCHK" in front, checksum behind
                if mtype != 0:
                    out.append((mtype, body, msize))
                p = body + msize
        return out

    # ---- data types
    def datatype(self, p: int) -> Tuple[_Type, int]:
        """Decoded type and the number of bytes the message occupies."""
        cls, ver = self.buf[p] & 0xF, self.buf[p] >> 4
        b0, b1 = self.buf[p + 1], self.buf[p + 2]
        size = self.u(p + 4, 4)
        q = p + 8
        order = ">" if b0 & 1 else "<"
        if cls == 0:
            return _Type(np.dtype(f"{order}{'i' if b0 & 8 else 'u'}{size}"), size), 12
        if cls == 1:
            return _Type(np.dtype(f"{order}f{size}"), size), 20
        if cls == 3:
            return _Type(np.dtype(f"S{size}"), size), 8
        if cls == 4:                                           # bit field
            return _Type(np.dtype(f"{order}u{size}"), size), 12
        if cls == 7:                                           # reference: opaque address
            return _Type(np.dtype(f"V{size}"), size), 8
        if cls == 5:                                           # opaque: tag padded to 8 bytes
            return _Type(np.dtype(f"V{size}"), size), 8 + ((b0 + 7) & ~7)
        if cls == 8:                                           # enum -> its base integer
            base, used = self.datatype(q)
            n = b0 | (b1 << 8)
            q += used
            for _ in range(n):
                e = self.buf.find(b"\0", q)
                q = q + ((e - q + 8) & ~7) if ver < 3 else e + 1
            return _Type(base.dtype, size), q + n * base.size - p
        if cls == 9:
            base, used = self.datatype(q)
            return _Type(None, size, "string" if (b0 & 0xF) == 1 else "sequence", base), 8 + used
        if cls == 6:
            n = b0 | (b1 << 8)
            names, formats, offsets = [], [], []
            for _ in range(n):
                e = self.buf.find(b"\0", q)
                names.append(self.buf[q:e].decode("utf8"))
                q = q + ((e - q + 8) & ~7) if ver < 3 else e + 1
                if ver == 1:
                    offsets.append(self.u(q, 4))
                    q += 4 + 1 + 3 + 4 + 4 + 16
                elif ver == 2:
                    offsets.append(self.u(q, 4))
                    q += 4
                else:
                    nb = 1 if size < 256 else 2 if size < 65536 else 3 if size < (1 << 24) else 4
                    offsets.append(self.u(q, nb))
                    q += nb
                member, used = self.datatype(q)
                if member.dtype is None:
                    raise Hdf5Unsupported("variable-length members of compound types")
                formats.append(member.dtype)
                q += used
            return _Type(np.dtype({"names": names, "formats": formats, "offsets": offsets, "itemsize": size}), size), q - p
        if cls == 10:
            rank = self.buf[q]
            q += 4 if ver < 3 else 1
            dims = [self.u(q + 4 * i, 4) for i in range(rank)]
            q += 4 * rank * (2 if ver < 3 else 1)
            base, used = self.datatype(q)
            if base.dtype is None:
                raise Hdf5Unsupported("arrays of variable-length elements")
            return _Type(np.dtype((base.dtype, tuple(dims))), size), q + used - p
        raise Hdf5Unsupported(f"datatype class {cls}")

    def dataspace(self, p: int) -> Tuple[int, ...]:
        ver, rank, flags = self.buf[p], self.buf[p + 1], self.buf[p + 2]
        if ver == 1:
            q = p + 8
        elif ver == 2:
            if self.buf[p + 3] == 2:                           # null dataspace
                return (0,)
            q = p + 4
        else:
            raise Hdf5Unsupported(f"dataspace version {ver}")
        return tuple(self.length(q + i * self.L) for i in range(rank))

    # ---- heaps
    def local_heap_data(self, heap: int) -> int:
        if self.buf[heap:heap + 4] != b"HEAP":
            raise Hdf5Error(f"{self.path}: no local heap at {heap:#x}")
        return self.addr(heap + 8 + 2 * self.L)

    def cstring(self, p: int) -> str:
        return self.buf[p:self.buf.find(b"\0", p)].decode("utf8", "replace")

    def global_heap_object(self, collection: int, index: int) -> bytes:
        table = self._gheaps.get(collection)
        if table is None:
            if self.buf[collection:collection + 4] != b"GCOL":
                raise Hdf5Error(f"{self.path}: no global heap collection at {collection:#x}")
            end = collection + self.length(collection + 8)
            p = collection + 8 + self.L
            table = {}
            while p + 8 + self.L <= end:
                idx, size = self.u(p, 2), self.length(p + 8)
                if idx == 0:
                    break
                table[idx] = (p + 8 + self.L, size)
                p += 8 + self.L + ((size + 7) & ~7)
            self._gheaps[collection] = table
        start, size = table[index]
        return self.buf[start:start + size]

    # ---- groups
    def symbol_table_links(self, btree: int, heap: int) -> Iterator[Tuple[str, int]]:
        data = self.local_heap_data(heap)
        entry = 2 * self.O + 24
        stack = [btree]
        while stack:
            node = stack.pop()
            sig = self.buf[node:node + 4]
            if sig == b"TREE":
                n = self.u(node + 6, 2)
                p = node + 8 + 2 * self.O + self.L             # first child (after key 0)
                kids = [self.addr(p + i * (self.O + self.L)) for i in range(n)]
                stack.extend(reversed(kids))
            elif sig == b"SNOD":
                n = self.u(node + 6, 2)
                for i in range(n):
                    p = node + 8 + i * entry
                    yield self.cstring(data + self.u(p, self.O)), self.addr(p + self.O)
            else:
                raise Hdf5Error(f"{self.path}: bad group node at {node:#x}")

    def link_message(self, p: int) -> Optional[Tuple[str, int]]:
        flags = self.buf[p + 1]
        q = p + 2
        ltype = 0
        if flags & 0x8:
            ltype = self.buf[q]
            q += 1
        if flags & 0x4:
            q += 8
        if flags & 0x10:
            q += 1
        nsz = 1 << (flags & 3)
        n = self.u(q, nsz)
        q += nsz
        name = self.buf[q:q + n].decode("utf8", "replace")
        return (name, self.addr(q + n)) if ltype == 0 else None        # soft / external links are skipped

    # ---- chunk index
    def chunk_btree(self, node: int, rank: int) -> Iterator[Tuple[Tuple[int, ...], int, int, int]]:
        """(chunk offset per dimension, address, stored size, filter mask) of every chunk."""
        if node == UNDEF:
            return
        if self.buf[node:node + 4] != b"TREE" or self.buf[node + 4] != 1:
            raise Hdf5Error(f"{self.path}: bad chunk index node at {node:#x}")
        level, n = self.buf[node + 5], self.u(node + 6, 2)
        key = 8 + 8 * (rank + 1)
        p = node + 8 + 2 * self.O
        for _ in range(n):
            size, mask = self.u(p, 4), self.u(p + 4, 4)
            offs = struct.unpack_from(f"<{rank}Q", self.buf, p + 8)
            child = self.addr(p + key)
            if level == 0:
                yield offs, child, size, mask
            else:
                yield from self.chunk_btree(child, rank)
            p += key + self.O


# ------------------------------------------------------------------------------------------------ objects
class _Object:
    def __init__(self, img: _Image, header: int, name: str):
        self._img, self._header, self.name = img, header, name
        self._msgs = img.messages(header)

    def _first(self, mtype: int) -> Optional[Tuple[int, int]]:
        for t, body, size in self._msgs:
            if t == mtype:
                return body, size
        return None

    @property
    def attrs(self) -> Dict[str, object]:
        img, out = self._img, {}
        for t, p, _ in self._msgs:
            if t != M_ATTRIBUTE:
                continue
            ver = img.buf[p]
            nlen, tlen, slen = img.u(p + 2, 2), img.u(p + 4, 2), img.u(p + 6, 2)
            q = p + 8 + (1 if ver == 3 else 0)
            pad = (lambda n: (n + 7) & ~7) if ver == 1 else (lambda n: n)
            name = img.buf[q:q + nlen].split(b"\0")[0].decode("utf8", "replace")
            q += pad(nlen)
            typ, _ = img.datatype(q)
            q += pad(tlen)
            shape = img.dataspace(q) if slen else ()
            q += pad(slen)
            count = int(np.prod(shape)) if shape else 1
            val = _decode(img, typ, img.buf[q:q + count * typ.size], count).reshape(shape)
            out[name] = val[()] if val.shape == () else val
        return out


def _decode(img: _Image, typ: _Type, raw: bytes, count: int) -> np.ndarray:
    if typ.vlen is None:
        return np.frombuffer(raw, dtype=typ.dtype, count=count).copy()
    out = np.empty(count, dtype=object)
    step = typ.size
    for i in range(count):
        p = i * step
        n = int.from_bytes(raw[p:p + 4], "little")
        a = int.from_bytes(raw[p + 4:p + 4 + img.O], "little")
        idx = int.from_bytes(raw[p + 4 + img.O:p + 8 + img.O], "little")
        if n == 0 or a == 0:
            blob = b""
        else:
            blob = img.global_heap_object(a + img.base, idx)
        if typ.vlen == "string":
            out[i] = blob[:n]
        else:
            out[i] = np.frombuffer(blob, dtype=typ.base.dtype, count=n).copy()
    return out


class Dataset(_Object):
    def __init__(self, img, header, name):
        super().__init__(img, header, name)
        sp, tp = self._first(M_DATASPACE), self._first(M_DATATYPE)
        if sp is None or tp is None or self._first(M_LAYOUT) is None:
            raise Hdf5Error(f"{name}: data set without dataspace / datatype / layout")
        self.shape = img.dataspace(sp[0])
        self._type, _ = img.datatype(tp[0])
        self.dtype = self._type.dtype if self._type.dtype is not None else np.dtype(object)
        self._chunk_index = None

    def __len__(self):
        return self.shape[0] if self.shape else 0

    @property
    def size(self):
        return int(np.prod(self.shape)) if self.shape else 1

    def _filters(self) -> List[Tuple[int, List[int]]]:
        m = self._first(M_FILTERS)
        if m is None:
            return []
        img, p = self._img, m[0]
        ver, n = img.buf[p], img.buf[p + 1]
        q = p + (8 if ver == 1 else 2)
        out = []
        for _ in range(n):
            fid = img.u(q, 2)
            q += 2
            nlen = 0
            if ver == 1 or fid >= 256:
                nlen = img.u(q, 2)
                q += 2
            q += 2                                              # flags
            nvals = img.u(q, 2)
            q += 2
            q += ((nlen + 7) & ~7) if ver == 1 else nlen
            vals = [img.u(q + 4 * i, 4) for i in range(nvals)]
            q += 4 * nvals
            if ver == 1 and nvals % 2:
                q += 4
            out.append((fid, vals))
        return out

    def _unfilter(self, raw: bytes, filters, mask: int) -> bytes:
        for k in range(len(filters) - 1, -1, -1):
            if mask & (1 << k):
                continue
            fid, vals = filters[k]
            if fid == 1:
                raw = zlib.decompress(raw)
            elif fid == 2:
                n = vals[0] if vals else self._type.size
                a = np.frombuffer(raw, dtype=np.uint8)
                whole = (a.size // n) * n
                raw = a[:whole].reshape(n, -1).T.tobytes() + a[whole:].tobytes()
            elif fid == 3:
                raw = raw[:-4]
            else:
                raise Hdf5Unsupported(f"{self.name}: filter {fid} (only deflate, shuffle and fletcher32 are built in)")
        return raw

    def _read_all(self, rows: Optional[Tuple[int, int]] = None) -> np.ndarray:
        """The whole data set, or rows ``[rows[0], rows[1])`` of its first axis (only those bytes / chunks are touched)."""
        img = self._img
        p, _ = self._first(M_LAYOUT)
        ver = img.buf[p]
        count, esize = self.size, self._type.size
        shape = self.shape
        if not shape:
            rows = None
        cut = (lambda a: a[rows[0]:rows[1]]) if rows else (lambda a: a)
        if ver in (1, 2):
            rank, cls = img.buf[p + 1], img.buf[p + 2]
            q = p + 8
            address = UNDEF
            if cls != 0:
                address = img.addr(q)
                q += img.O
            dims = [img.u(q + 4 * i, 4) for i in range(rank)]
            q += 4 * rank
            if cls == 0:
                raw = img.buf[q + 4:q + 4 + img.u(q, 4)]
                return cut(_decode(img, self._type, raw, count).reshape(shape))
            if cls == 1:
                return self._contiguous(address, count, esize, rows)
            return self._chunked(address, dims[:-1], None, rows)
        cls = img.buf[p + 1]
        if cls == 0:
            n = img.u(p + 2, 2)
            return cut(_decode(img, self._type, img.buf[p + 4:p + 4 + n], count).reshape(shape))
        if cls == 1:
            return self._contiguous(img.addr(p + 2), count, esize, rows)
        if cls != 2:
            raise Hdf5Unsupported(f"{self.name}: layout class {cls}")
        if ver == 3:
            rank = img.buf[p + 2]
            address = img.addr(p + 3)
            dims = [img.u(p + 3 + img.O + 4 * i, 4) for i in range(rank)]
            return self._chunked(address, dims[:-1], None, rows)
        # version 4: flags, rank, size of a dimension field, dims, index type
        flags, rank, dsz = img.buf[p + 2], img.buf[p + 3], img.buf[p + 4]
        dims = [img.u(p + 5 + dsz * i, dsz) for i in range(rank)]
        q = p + 5 + dsz * rank
        itype = img.buf[q]
        q += 1
        if itype == 1:                                          # single chunk
            size, mask = None, 0
            if flags & 2:
                size, mask = img.length(q), img.u(q + img.L, 4)
                q += img.L + 4
            return self._chunked(UNDEF, dims[:-1], [(tuple([0] * (rank - 1)), img.addr(q), size, mask)], rows)
        if itype == 2:                                          # implicit: chunks back to back, no filters
            base = img.addr(q)
            cdims = dims[:-1]
            per = int(np.prod(cdims)) * esize
            grid = [-(-s // c) for s, c in zip(shape, cdims)]
            chunks = []
            for k, idx in enumerate(np.ndindex(*grid)):
                chunks.append((tuple(i * c for i, c in zip(idx, cdims)), base + k * per, per, 0))
            return self._chunked(UNDEF, cdims, chunks, rows)
        raise Hdf5Unsupported(f"{self.name}: chunk index type {itype} (written with libver='latest'); re-save the file with "
                              "the default library version bounds")

    def _contiguous(self, address, count, esize, rows=None):
        shape = self.shape
        if rows:
            per_row = int(np.prod(shape[1:])) if len(shape) > 1 else 1
            shape = (rows[1] - rows[0],) + tuple(shape[1:])
            count = shape[0] * per_row
            if address != UNDEF:
                address += rows[0] * per_row * esize
        if address == UNDEF:                                    # never written: fill value (zeros)
            if self._type.vlen:
                out = np.empty(count, dtype=object)
                out[:] = b""
                return out.reshape(shape)
            return np.zeros(shape, dtype=self._type.dtype)
        raw = self._img.buf[address:address + count * esize]
        return _decode(self._img, self._type, raw, count).reshape(shape)

    def _chunked(self, btree, cdims, chunks, rows=None) -> np.ndarray:
        img, typ = self._img, self._type
        rank = len(cdims)
        filters = self._filters()
        lo, hi = rows if rows else (0, self.shape[0] if self.shape else 1)
        shape = ((hi - lo,) + tuple(self.shape[1:])) if self.shape else ()
        if typ.vlen:
            out = np.empty(shape, dtype=object)
            out.fill(b"")
        else:
            out = np.zeros(shape, dtype=typ.dtype)
        if chunks is None:
            if self._chunk_index is None:                       # the B-tree walk is the expensive part: keep the flat list
                self._chunk_index = list(img.chunk_btree(btree, rank))
            chunks = self._chunk_index
        per = int(np.prod(cdims))
        for offs, address, size, mask in chunks:
            if address == UNDEF or (rank and (offs[0] + cdims[0] <= lo or offs[0] >= hi)):
                continue
            raw = img.buf[address:address + (size if size is not None else per * typ.size)]
            if filters:
                raw = self._unfilter(raw, filters, mask)
            block = _decode(img, typ, raw, per).reshape(cdims)
            if not rank:
                out[()] = block[()]
                continue
            # intersection of the chunk with [lo, hi) x the data set's extent, in data set and in chunk coordinates
            a0, b0 = max(offs[0], lo), min(offs[0] + cdims[0], hi, self.shape[0])
            dst = (slice(a0 - lo, b0 - lo),) + tuple(slice(o, min(o + c, s_)) for o, c, s_ in zip(offs[1:], cdims[1:], self.shape[1:]))
            src = (slice(a0 - offs[0], b0 - offs[0]),) + tuple(slice(0, d.stop - d.start) for d in dst[1:])
            out[dst] = block[src]
        return out

    def __getitem__(self, key):
        """h5py-style indexing.  An integer or a slice on the first axis reads only those rows from the file."""
        first, rest = (key[0], key[1:]) if isinstance(key, tuple) and key else (key, ())
        n = self.shape[0] if self.shape else 0
        if self.shape and isinstance(first, (int, np.integer)) and not isinstance(first, bool):
            i = int(first) + (n if first < 0 else 0)
            if not 0 <= i < n:
                raise IndexError(f"index {int(first)} out of range for axis 0 with size {n}")
            row = self._read_all((i, i + 1))[0]
            return row[rest] if rest else row
        if self.shape and isinstance(first, slice):
            lo, hi, step = first.indices(n)
            if step > 0 and hi > lo:
                part = self._read_all((lo, hi))[::step]
                return part[(slice(None),) + rest] if rest else part
        data = self._read_all()
        if key is Ellipsis or (isinstance(key, tuple) and len(key) == 0):
            return data[()] if data.shape == () else data
        return data[key]

    def __array__(self, dtype=None, copy=None):
        a = self._read_all()
        return a.astype(dtype) if dtype is not None else a

    def asstr(self, encoding="utf8"):
        return np.array([b.decode(encoding) if isinstance(b, bytes) else b for b in self._read_all().reshape(-1)],
                        dtype=object).reshape(self.shape)

    def __repr__(self):
        return f'<minihdf5 dataset "{self.name}": shape {self.shape}, type {self.dtype}>'


class Group(_Object):
    def __init__(self, img, header, name):
        super().__init__(img, header, name)
        self._links: Optional[Dict[str, int]] = None

    def _load(self) -> Dict[str, int]:
        if self._links is None:
            img, links = self._img, {}
            st = self._first(M_SYMTAB)
            if st is not None:
                links.update(img.symbol_table_links(img.addr(st[0]), img.addr(st[0] + img.O)))
            for t, p, _ in self._msgs:
                if t == M_LINK:
                    ln = img.link_message(p)
                    if ln:
                        links[ln[0]] = ln[1]
                elif t == M_LINK_INFO:
                    flags = img.buf[p + 1]
                    q = p + 2 + (8 if flags & 1 else 0)
                    if img.addr(q) != UNDEF:
                        raise Hdf5Unsupported(f"{self.name or '/'}: densely stored group (file written with libver='latest'); "
                                              "re-save it with the default library version bounds")
            self._links = links
        return self._links

    def keys(self):
        return list(self._load().keys())

    def __iter__(self):
        return iter(self.keys())

    def __len__(self):
        return len(self._load())

    def __contains__(self, name) -> bool:
        try:
            self[name]
            return True
        except KeyError:
            return False

    def _open(self, name: str, header: int):
        path = f"{self.name.rstrip('/')}/{name}"
        types = {t for t, _, _ in self._img.messages(header)}
        if M_LAYOUT in types and M_DATATYPE in types:
            return Dataset(self._img, header, path)
        return Group(self._img, header, path)

    def __getitem__(self, name: str):
        node = self
        for part in [p for p in str(name).split("/") if p]:
            if not isinstance(node, Group):
                raise KeyError(name)
            links = node._load()
            if part not in links:
                raise KeyError(f"{name!r} (no {part!r} in {node.name or '/'})")
            node = node._open(part, links[part])
        return node

    def get(self, name, default=None):
        try:
            return self[name]
        except KeyError:
            return default

    def items(self):
        return [(k, self[k]) for k in self.keys()]

    def values(self):
        return [self[k] for k in self.keys()]

    def visit(self, fn, _prefix=""):
        for k in self.keys():
            child = self[k]
            path = f"{_prefix}{k}"
            r = fn(path)
            if r is not None:
                return r
            if isinstance(child, Group):
                r = child.visit(fn, path + "/")
                if r is not None:
                    return r
        return None

    def __repr__(self):
        return f'<minihdf5 group "{self.name or "/"}" ({len(self)} members)>'


class File(Group):
    """Read-only. ``mode`` is accepted for h5py compatibility and must be ``"r"``."""

    def __init__(self, path: str, mode: str = "r"):
        if mode != "r":
            raise Hdf5Unsupported("minihdf5.File is read-only; use write_hdf5(path, tree) to create files")
        img = _Image(path)
        try:
            super().__init__(img, img.root_header, "")
        except Exception:
            img.close()
            raise
        self.filename = path

    def close(self):
        self._img.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False


# ------------------------------------------------------------------------------------------------ writing
class H5Data:
    """A data set with storage options: ``chunks`` (tuple) switches to chunked storage, ``compression`` is a deflate level."""

    def __init__(self, data, chunks: Optional[Tuple[int, ...]] = None, compression: Optional[int] = None, shuffle: bool = False):
        self.data, self.chunks, self.compression, self.shuffle = data, chunks, compression, shuffle
        if (compression is not None or shuffle) and chunks is None:
            raise ValueError("filters need chunked storage: pass chunks=(...)")


_LEAF_K, _INTERNAL_K, _CHUNK_K = 4, 16, 32
_O = _L = 8


def _pad8(b: bytes) -> bytes:
    return b + b"\0" * (-len(b) % 8)


def _encode_type(dt: np.dtype) -> bytes:
    dt = np.dtype(dt)
    if dt.kind in "iu":
        return struct.pack("<BBBBIHH", 0x10, 0x08 if dt.kind == "i" else 0, 0, 0, dt.itemsize, 0, 8 * dt.itemsize)
    if dt.kind == "f":
        exp_size, mant = {2: (5, 10), 4: (8, 23), 8: (11, 52)}[dt.itemsize]
        return struct.pack("<BBBBIHHBBBBI", 0x11, 0x20, 8 * dt.itemsize - 1, 0, dt.itemsize, 0, 8 * dt.itemsize,
                           mant, exp_size, 0, mant, (1 << (exp_size - 1)) - 1)
    if dt.kind == "S":
        return struct.pack("<BBBBI", 0x13, 0x01, 0, 0, dt.itemsize)
    raise Hdf5Unsupported(f"cannot store numpy dtype {dt}")


_VLEN_STR = struct.pack("<BBBBI", 0x19, 0x01, 0x01, 0, 4 + _O + 4) + struct.pack("<BBBBI", 0x13, 0x10, 0, 0, 1)


class _Writer:
    def __init__(self):
        self.buf = bytearray(96)                                # superblock written last

    def alloc(self, data: bytes) -> int:
        self.buf.extend(b"\0" * (-len(self.buf) % 8))
        at = len(self.buf)
        self.buf.extend(data)
        return at

    # ---- object header, version 1
    def header(self, msgs: List[Tuple[int, bytes]]) -> int:
        body = b"".join(struct.pack("<HHBBH", t, len(_pad8(m)), 0, 0, 0) + _pad8(m) for t, m in msgs)
        return self.alloc(struct.pack("<BBHII", 1, 0, len(msgs), 1, len(body)) + b"\0" * 4 + body)

    # ---- variable-length strings: global heap collections
    def vlen_strings(self, items: List[bytes]) -> bytes:
        refs = bytearray()
        i = 0
        while i < len(items):
            objs = bytearray()
            first = i
            while i < len(items) and i - first < 65000 and len(objs) < (1 << 20):
                b = items[i]
                objs += struct.pack("<HHIQ", i - first + 1, 1, 0, len(b)) + _pad8(b)
                i += 1
            size = max(4096, 16 + len(objs) + 16)
            size += -size % 8
            free = size - 16 - len(objs)
            blob = b"GCOL" + struct.pack("<BBBBQ", 1, 0, 0, 0, size) + bytes(objs) + struct.pack("<HHIQ", 0, 0, 0, free)
            at = self.alloc(blob + b"\0" * (size - len(blob)))
            for k in range(first, i):
                refs += struct.pack("<IQI", len(items[k]), at, k - first + 1)
        return bytes(refs)

    # ---- data sets
    def dataset(self, spec) -> int:
        opts = spec if isinstance(spec, H5Data) else H5Data(spec)
        data = opts.data
        strings = None
        if isinstance(data, (list, tuple)) and data and isinstance(data[0], (str, bytes)):
            data = np.array(list(data), dtype=object)
        arr = np.asarray(data)
        if arr.dtype == object or arr.dtype.kind == "U":
            strings = [s.encode("utf8") if isinstance(s, str) else bytes(s) for s in arr.reshape(-1)]
        elif arr.dtype == np.bool_:
            arr = arr.astype(np.uint8)
        if arr.dtype.byteorder == ">":
            arr = arr.astype(arr.dtype.newbyteorder("<"))
        shape = arr.shape
        if strings is not None:
            tmsg, esize = _VLEN_STR, 4 + _O + 4
            raw_all = self.vlen_strings(strings)
            cells = np.frombuffer(raw_all, dtype=f"V{esize}").reshape(shape)
        else:
            tmsg, esize = _encode_type(arr.dtype), arr.dtype.itemsize
            cells = np.ascontiguousarray(arr)
        space = struct.pack("<BBBBI", 1, len(shape), 0, 0, 0) + b"".join(struct.pack("<Q", d) for d in shape)
        msgs = [(M_DATASPACE, space), (M_DATATYPE, tmsg)]
        if opts.chunks is None:
            raw = cells.tobytes()
            at = self.alloc(raw) if raw else UNDEF
            msgs += [(M_FILL, struct.pack("<BBBB", 2, 2, 2, 0)),
                     (M_LAYOUT, struct.pack("<BBQQ", 3, 1, at, len(raw)))]
            return self.header(msgs)
        cdims = tuple(int(c) for c in opts.chunks)
        if len(cdims) != len(shape) or any(c <= 0 for c in cdims):
            raise ValueError(f"chunks {cdims} do not fit shape {shape}")
        filters = []
        if opts.shuffle:
            filters.append((2, [esize]))
        if opts.compression is not None:
            filters.append((1, [int(opts.compression)]))
        entries = []
        grid = [-(-s // c) for s, c in zip(shape, cdims)]
        for idx in np.ndindex(*grid):
            offs = tuple(i * c for i, c in zip(idx, cdims))
            block = np.zeros(cdims, dtype=cells.dtype)
            sel = tuple(slice(o, min(o + c, s)) for o, c, s in zip(offs, cdims, shape))
            part = cells[sel]
            block[tuple(slice(0, n) for n in part.shape)] = part
            raw = block.tobytes()
            for fid, vals in filters:
                if fid == 2:
                    a = np.frombuffer(raw, dtype=np.uint8)
                    raw = a.reshape(-1, vals[0]).T.tobytes()
                else:
                    raw = zlib.compress(raw, vals[0])
            entries.append((offs, self.alloc(raw), len(raw)))
        btree = self.chunk_index(entries, shape, cdims, esize)
        msgs.append((M_FILL, struct.pack("<BBBB", 2, 3, 2, 0)))
        msgs.append((M_LAYOUT, struct.pack("<BBBQ", 3, 2, len(shape) + 1, btree) +
                     b"".join(struct.pack("<I", c) for c in cdims + (esize,))))
        if filters:
            f = struct.pack("<BBHI", 1, len(filters), 0, 0)
            for fid, vals in filters:
                f += struct.pack("<HHHH", fid, 0, 1, len(vals)) + b"".join(struct.pack("<I", v) for v in vals)
                if len(vals) % 2:
                    f += b"\0" * 4
            msgs.append((M_FILTERS, f))
        return self.header(msgs)

    def chunk_index(self, entries, shape, cdims, esize) -> int:
        rank = len(shape)
        key = lambda size, offs: struct.pack("<II", size, 0) + b"".join(struct.pack("<Q", o) for o in offs) + struct.pack("<Q", 0)  # noqa: E731
        end_offs = tuple(-(-s // c) * c for s, c in zip(shape, cdims))
        # level 0 nodes hold the chunks, higher levels hold nodes; (first key, address) per child
        level, kids = 0, [(key(sz, offs), at) for offs, at, sz in entries]
        cap = 2 * _CHUNK_K
        node_bytes = 24 + cap * (len(key(0, end_offs)) + _O) + len(key(0, end_offs))
        while True:
            groups = [kids[i:i + cap] for i in range(0, len(kids), cap)] or [[]]
            addrs = [self.alloc(b"\0" * node_bytes) for _ in groups]
            for gi, (g, at) in enumerate(zip(groups, addrs)):
                left = addrs[gi - 1] if gi > 0 else UNDEF
                right = addrs[gi + 1] if gi + 1 < len(addrs) else UNDEF
                body = b"TREE" + struct.pack("<BBHQQ", 1, level, len(g), left, right)
                for k, a in g:
                    body += k + struct.pack("<Q", a)
                last = groups[gi + 1][0][0] if gi + 1 < len(groups) else key(0, end_offs)
                body += last
                self.buf[at:at + len(body)] = body
            if len(groups) == 1:
                return addrs[0]
            kids = [(g[0][0], at) for g, at in zip(groups, addrs)]
            level += 1

    # ---- groups: local heap + symbol nodes + B-tree
    def group(self, tree: dict) -> Tuple[int, int, int]:
        children = []
        for name in sorted(tree, key=lambda s: str(s).encode("utf8")):
            v = tree[name]
            children.append((str(name).encode("utf8"), self.group(v)[0] if isinstance(v, dict) else self.dataset(v)))
        heap_data = bytearray(b"\0" * 8)
        offsets = []
        for nm, _ in children:
            offsets.append(len(heap_data))
            heap_data += _pad8(nm + b"\0")
        data_at = self.alloc(bytes(heap_data))
        heap_at = self.alloc(b"HEAP" + struct.pack("<BBBBQQQ", 0, 0, 0, 0, len(heap_data), 1, data_at))
        entry_bytes, cap = 2 * _O + 24, 2 * _LEAF_K
        leaves = []                                             # (offset of the largest name, address)
        for i in range(0, len(children), cap):
            part = list(zip(offsets[i:i + cap], children[i:i + cap]))
            body = b"SNOD" + struct.pack("<BBH", 1, 0, len(part))
            for off, (_, header) in part:
                body += struct.pack("<QQII", off, header, 0, 0) + b"\0" * 16
            at = self.alloc(body + b"\0" * (8 + cap * entry_bytes - len(body)))
            leaves.append((part[-1][0], at))
        level, kids, cap = 0, leaves, 2 * _INTERNAL_K
        node_bytes = 24 + cap * (_L + _O) + _L
        while True:
            groups = [kids[i:i + cap] for i in range(0, len(kids), cap)] or [[]]
            addrs = [self.alloc(b"\0" * node_bytes) for _ in groups]
            for gi, (g, at) in enumerate(zip(groups, addrs)):
                left = addrs[gi - 1] if gi > 0 else UNDEF
                right = addrs[gi + 1] if gi + 1 < len(addrs) else UNDEF
                body = b"TREE" + struct.pack("<BBHQQ", 0, level, len(g), left, right)
                body += struct.pack("<Q", groups[gi - 1][-1][0] if gi > 0 else 0)
                for k, a in g:
                    body += struct.pack("<QQ", a, k)
                self.buf[at:at + len(body)] = body
            if len(groups) == 1:
                btree = addrs[0]
                break
            kids = [(g[-1][0], at) for g, at in zip(groups, addrs)]
            level += 1
        header = self.header([(M_SYMTAB, struct.pack("<QQ", btree, heap_at))])
        return header, btree, heap_at

    def finish(self, root: Tuple[int, int, int]) -> bytes:
        header, btree, heap = root
        self.buf.extend(b"\0" * (-len(self.buf) % 8))
        sb = SIGNATURE + struct.pack("<BBBBBBBBHHI", 0, 0, 0, 0, 0, _O, _L, 0, _LEAF_K, _INTERNAL_K, 0)
        sb += struct.pack("<QQQQ", 0, UNDEF, len(self.buf), UNDEF)
        sb += struct.pack("<QQII", 0, header, 1, 0) + struct.pack("<QQ", btree, heap)
        assert len(sb) == 96
        self.buf[0:96] = sb
        return bytes(self.buf)


def write_hdf5(path: str, tree: dict) -> str:
    """``tree``: nested dicts (groups) whose leaves are numpy arrays, lists of ``str`` / ``bytes`` (variable-length strings)
    or :class:`H5Data` wrappers carrying chunking / compression options."""
    w = _Writer()
    blob = w.finish(w.group(tree))
    with open(path, "wb") as fh:
        fh.write(blob)
    return path
