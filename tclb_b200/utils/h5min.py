"""Reader for the HDF5 subset written by csrc/io/h5min.cpp (and by libhdf5 for plain contiguous datasets in a
version-0/1 file): superblock v0, v1 object headers, group B-tree + local heap + symbol-table nodes, contiguous layout or chunked layout (version-1 chunk B-tree) with the deflate filter,
little-endian IEEE floats and integers.  Follows the structures, i.e. it fails on an inconsistent file; used by the
tests and handy where h5py is missing.  ``read(path) -> {name: ndarray}``."""
from __future__ import annotations

import struct
from typing import Dict

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF


def _u(b, off, n):
    return int.from_bytes(b[off:off + n], "little")


def _messages(b, addr):
    ver, _, nmsg, refcnt, size = struct.unpack_from("<BBHII", b, addr)
    assert ver == 1, "object header version"
    off, end, out = addr + 16, addr + 16 + size, []
    for _ in range(nmsg):
        mtype, msize, flags = struct.unpack_from("<HHB", b, off)
        out.append((mtype, b[off + 8:off + 8 + msize]))
        off += 8 + msize
        assert off <= end, "message overruns the object header"
    return out


def _dtype(msg):
    cls, ver = msg[0] & 15, msg[0] >> 4
    size = _u(msg, 4, 4)
    assert ver == 1 and (msg[1] & 1) == 0, "datatype version / byte order"
    if cls == 1:
        prec, eloc, esize, mloc, msize, bias = struct.unpack_from("<xxHBBBBI", msg, 8)
        assert (size, prec, esize, msize, bias) in ((4, 32, 8, 23, 127), (8, 64, 11, 52, 1023)), "not an IEEE float"
        return np.dtype("<f4" if size == 4 else "<f8")
    if cls == 0:
        signed = bool(msg[1] & 8)
        return np.dtype(("<i" if signed else "<u") + str(size))
    raise ValueError(f"datatype class {cls} not supported")


def read(path: str) -> Dict[str, np.ndarray]:
    b = open(path, "rb").read()
    assert b[:8] == b"\x89HDF\r\n\x1a\n", "signature"
    assert b[8] == 0 and b[13] == 8 and b[14] == 8, "superblock version 0 with 8-byte offsets expected"
    leaf_k, int_k = struct.unpack_from("<HH", b, 16)
    eof = _u(b, 40, 8)
    assert eof == len(b), f"end-of-file address {eof} != file size {len(b)}"
    root_hdr, cache = _u(b, 64, 8), _u(b, 72, 4)
    btree = heap = None
    for t, m in _messages(b, root_hdr):
        if t == 0x11:
            btree, heap = _u(m, 0, 8), _u(m, 8, 8)
    assert btree is not None, "root group has no symbol table"
    if cache == 1:
        assert (_u(b, 80, 8), _u(b, 88, 8)) == (btree, heap), "cached root addresses differ from the header message"
    assert b[heap:heap + 4] == b"HEAP"
    hsize, hfree, hdata = _u(b, heap + 8, 8), _u(b, heap + 16, 8), _u(b, heap + 24, 8)
    assert hfree == 1 or hfree == UNDEF or hfree + 16 <= hsize, "heap free list"

    def name_at(off):
        s = hdata + off
        return b[s:b.index(b"\0", s)].decode()

    out: Dict[str, np.ndarray] = {}

    def walk(node):
        assert b[node:node + 4] == b"TREE" and b[node + 4] == 0, "group B-tree node"
        level, used = b[node + 5], _u(b, node + 6, 2)
        p = node + 24
        for i in range(used):
            child = _u(b, p + 8, 8)
            if level > 0:
                walk(child)
            else:
                assert b[child:child + 4] == b"SNOD" and b[child + 4] == 1
                n = _u(b, child + 6, 2)
                assert n <= 2 * leaf_k
                names = []
                for e in range(n):
                    eo = child + 8 + 40 * e
                    nm, hdr = name_at(_u(b, eo, 8)), _u(b, eo + 8, 8)
                    names.append(nm)
                    out[nm] = _dataset(hdr)
                assert names == sorted(names), "symbol table entries are not ordered"
                assert name_at(_u(b, p + 16, 8)) == names[-1], "B-tree key is not the last name of its child"
            p += 16

    def _chunks(node, rank, out_list):
        """version-1 B-tree of raw data chunks: (offsets, address, stored size, filter mask) of every chunk, in key order"""
        assert b[node:node + 4] == b"TREE" and b[node + 4] == 1, "chunk B-tree node"
        level, used = b[node + 5], _u(b, node + 6, 2)
        assert used <= 64, "more entries than a node of K = 32 holds"
        ksz = 8 + 8 * (rank + 1)
        p = node + 24
        for i in range(used):
            size, mask = _u(b, p, 4), _u(b, p + 4, 4)
            offs = [_u(b, p + 8 + 8 * d, 8) for d in range(rank + 1)]
            child = _u(b, p + ksz, 8)
            if level > 0:
                _chunks(child, rank, out_list)
            else:
                out_list.append((offs, child, size, mask))
            p += ksz + 8

    def _dataset(hdr):
        import zlib
        dims = dt = addr = nbytes = chunk = None
        filters = []
        for t, m in _messages(b, hdr):
            if t == 1:
                assert m[0] == 1
                dims = [_u(m, 8 + 8 * i, 8) for i in range(m[1])]
            elif t == 3:
                dt = _dtype(m)
            elif t == 0x0B:
                assert m[0] == 1, "filter pipeline version 1 expected"
                off = 8
                for _ in range(m[1]):
                    fid, nlen, flags, ncl = struct.unpack_from("<HHHH", m, off)
                    off += 8 + ((nlen + 7) // 8) * 8
                    vals = [_u(m, off + 4 * i, 4) for i in range(ncl)]
                    off += 4 * ncl + (4 if ncl % 2 else 0)
                    filters.append((fid, vals))
            elif t == 8:
                assert m[0] == 3, "layout version 3 expected"
                if m[1] == 1:
                    addr, nbytes = _u(m, 2, 8), _u(m, 10, 8)
                elif m[1] == 2:
                    nd = m[2]
                    addr = _u(m, 3, 8)
                    chunk = [_u(m, 11 + 4 * i, 4) for i in range(nd)]
                else:
                    raise ValueError("layout class %d not supported" % m[1])
        assert dims is not None and dt is not None and addr is not None, "dataset header incomplete"
        n = int(np.prod(dims)) if dims else 1
        if chunk is None:
            assert nbytes == n * dt.itemsize and addr + nbytes <= len(b), "data extent"
            return np.frombuffer(b, dtype=dt, count=n, offset=addr).reshape(dims).copy()
        rank = len(dims)
        assert len(chunk) == rank + 1 and chunk[-1] == dt.itemsize, "chunk dimensions"
        assert all(f[0] == 1 for f in filters), "only the deflate filter is supported"
        out_arr = np.zeros(dims, dtype=dt)
        lst = []
        _chunks(addr, rank, lst)
        assert [c[0] for c in lst] == sorted(c[0] for c in lst), "chunk keys are not ordered"
        for offs, caddr, size, mask in lst:
            raw = b[caddr:caddr + size]
            assert len(raw) == size, "chunk beyond the end of the file"
            if filters and not (mask & 1):
                raw = zlib.decompress(raw)
            shape = chunk[:-1]
            assert len(raw) == int(np.prod(shape)) * dt.itemsize, "chunk size after the filters"
            blk = np.frombuffer(raw, dtype=dt).reshape(shape)
            sl_out = tuple(slice(offs[d], min(offs[d] + shape[d], dims[d])) for d in range(rank))
            sl_in = tuple(slice(0, s_.stop - s_.start) for s_ in sl_out)
            out_arr[sl_out] = blk[sl_in]
        covered = sum(min(c[0][0] + chunk[0], dims[0]) - c[0][0] for c in lst)
        assert covered == dims[0], "the chunks do not cover the dataset"
        return out_arr

    walk(btree)
    return out
