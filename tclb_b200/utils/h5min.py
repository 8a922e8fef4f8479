"""Reader for the HDF5 subset written by csrc/io/h5min.cpp (and by libhdf5 for plain contiguous datasets in a
version-0/1 file): superblock v0, v1 object headers, group B-tree + local heap + symbol-table nodes, contiguous layout,
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

    def _dataset(hdr):
        dims = dt = addr = nbytes = None
        for t, m in _messages(b, hdr):
            if t == 1:
                assert m[0] == 1
                dims = [_u(m, 8 + 8 * i, 8) for i in range(m[1])]
            elif t == 3:
                dt = _dtype(m)
            elif t == 8:
                assert m[0] == 3 and m[1] == 1, "contiguous layout version 3 expected"
                addr, nbytes = _u(m, 2, 8), _u(m, 10, 8)
        assert dims is not None and dt is not None and addr is not None, "dataset header incomplete"
        n = int(np.prod(dims)) if dims else 1
        assert nbytes == n * dt.itemsize and addr + nbytes <= len(b), "data extent"
        return np.frombuffer(b, dtype=dt, count=n, offset=addr).reshape(dims).copy()

    walk(btree)
    return out
