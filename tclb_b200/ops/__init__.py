"""Stand-alone device operations of the runtime on torch tensors (the kernels live in ``csrc/kernels``; a call here is a
ctypes call into ``libtclb_core.so`` on the tensor's device and the current torch stream — no copy, no fallback).

``scan``                 NaN count / min / max of an array (the Failcheck kernel; reference src/Handlers/cbFailcheck.cpp:40-74)
``particle_grid``        uniform-grid bucket table over particle rows (reference src/SolidGrid.hpp:11-60)
``accumulate_nonzero``   dst += src where src != 0 (the adjoint ghost fold's add; reference src/Lattice.cu.Rt:542-613)
"""
from __future__ import annotations

import ctypes as C
from typing import Tuple

from ..runtime import core

__all__ = ["scan", "particle_grid", "accumulate_nonzero"]


def _stream(t):
    import torch
    return C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream) if t.is_cuda else None


def _real_size(t) -> int:
    import torch
    if t.dtype == torch.float32:
        return 4
    if t.dtype == torch.float64:
        return 8
    raise TypeError("float32 or float64 tensor expected, got %s" % t.dtype)


def scan(t, static: bool = False):
    """(nan_count, min, max) of a contiguous CUDA tensor as a float64 device tensor of 3.  ``static=True`` uses the entry
    that allocates nothing (fixed scratch in the library; arrays up to 2^26 elements)."""
    import torch
    if not (t.is_cuda and t.is_contiguous()):
        raise ValueError("scan needs a contiguous CUDA tensor")
    lib = core()
    out = torch.empty(3, dtype=torch.float64, device=t.device)
    with torch.cuda.device(t.device):
        if static:
            lib.tclb_scan_array_static.argtypes = [C.c_void_p, C.c_longlong, C.c_int, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int]
            err = C.create_string_buffer(256)
            if lib.tclb_scan_array_static(t.data_ptr(), t.numel(), _real_size(t), _stream(t), out.data_ptr(), err, len(err)) != 0:
                raise RuntimeError(err.value.decode())
        else:
            lib.tclb_scan_array.argtypes = [C.c_void_p, C.c_longlong, C.c_int, C.c_void_p, C.c_void_p]
            lib.tclb_scan_last_error.restype = C.c_char_p
            if lib.tclb_scan_array(t.data_ptr(), t.numel(), _real_size(t), _stream(t), out.data_ptr()) != 0:
                raise RuntimeError(lib.tclb_scan_last_error().decode())
    return out


def particle_grid(rows, origin, delta: float, cells: Tuple[int, int, int], depth: int = 8):
    """Bucket table of particles: ``rows`` is [n, 16] (radius, position, ...; column 0 = r, 1..3 = position).  Returns
    (grid int32[cz, cy, cx, depth] with -1 for empty slots, counts int32[cz, cy, cx], overflow int32[1]); a particle is
    entered in the cell of its centre (cell edge >= 2 * largest reach, so a node looks at the 8 cells around it)."""
    import torch
    if rows.dim() != 2 or rows.shape[1] != 16 or not rows.is_contiguous():
        raise ValueError("rows must be a contiguous [n, 16] tensor")
    cx, cy, cz = (int(v) for v in cells)
    dev = rows.device
    grid = torch.empty((cz, cy, cx, depth), dtype=torch.int32, device=dev)
    counts = torch.empty((cz, cy, cx), dtype=torch.int32, device=dev)
    overflow = torch.empty(1, dtype=torch.int32, device=dev)
    lib = core()
    lib.tclb_particle_grid_fill.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_double), C.c_double, C.POINTER(C.c_int), C.c_int,
                                            C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    org = (C.c_double * 3)(*[float(v) for v in origin])
    gn = (C.c_int * 3)(cx, cy, cz)
    ctx = torch.cuda.device(dev) if rows.is_cuda else _Null()
    with ctx:
        rc = lib.tclb_particle_grid_fill(rows.data_ptr(), rows.shape[0], _real_size(rows), org, 1.0 / float(delta), gn, depth,
                                         grid.data_ptr(), counts.data_ptr(), overflow.data_ptr(), 1 if rows.is_cuda else 0, _stream(rows))
    if rc != 0:
        raise RuntimeError("tclb_particle_grid_fill failed")
    return grid, counts, overflow


def accumulate_nonzero(dst, src):
    """dst[i] += src[i] for the elements where src[i] != 0, in place (contiguous tensors of one dtype and device)."""
    import torch
    if dst.shape != src.shape or dst.dtype != src.dtype or dst.device != src.device or not (dst.is_contiguous() and src.is_contiguous()):
        raise ValueError("dst and src must be contiguous tensors of one shape, dtype and device")
    lib = core()
    lib.tclb_adjoint_add_into.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_int, C.c_void_p]
    ctx = torch.cuda.device(dst.device) if dst.is_cuda else _Null()
    with ctx:
        if lib.tclb_adjoint_add_into(dst.data_ptr(), src.data_ptr(), dst.numel(), _real_size(dst), 1 if dst.is_cuda else 0, _stream(dst)) != 0:
            raise RuntimeError("tclb_adjoint_add_into failed")
    return dst


class _Null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
