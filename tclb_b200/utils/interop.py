"""Zero-copy views of lattice memory as torch tensors (the Python coupling surface: the reference's
<RunPython> hands numpy COPIES through reticulate, src/Handlers/cbRunR.cpp:104-150; here fields are
views of the device slab)."""
from __future__ import annotations

import numpy as np


class _CudaArray:
    """Minimal object speaking __cuda_array_interface__ so torch.as_tensor() wraps a raw device pointer."""

    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(int(s) for s in shape), "typestr": typestr,
                                         "data": (int(ptr), False), "version": 2, "strides": None}


def tensor_from_ptr(ptr: int, shape, dtype, device: int):
    import torch
    ts = {torch.float32: "<f4", torch.float64: "<f8", torch.float16: "<f2", torch.uint16: "<u2", torch.int16: "<i2", torch.int32: "<i4", torch.uint8: "|u1"}[dtype]
    return torch.as_tensor(_CudaArray(ptr, shape, ts), device=torch.device("cuda", device))


def snapshot_tensor(domain, rank: int = None, snap: int = -1):
    """[n_fields][nz+2gz][ny+2gy][nx] view of a rank's snapshot (current one by default).  The element type is the
    STORAGE type of the library (float16 for *_half builds) and, for *shift builds, the values are deviations from
    the per-field shift (Model.field_shift)."""
    import torch
    rank = domain.local[0] if rank is None else rank
    ptr, shp, gy, gz, cur = domain.snapshot_info(rank, snap)
    dtype = {2: torch.float16, 4: torch.float32, 8: torch.float64}[int(domain.model.d.storage_size)]
    if domain.backend == 1:
        dev = domain.rank_device(rank) if hasattr(domain, "rank_device") else torch.cuda.current_device()
        return tensor_from_ptr(ptr, shp, dtype, dev), gy, gz
    import ctypes as C
    n = int(np.prod(shp)) * int(domain.model.d.storage_size)
    buf = (C.c_char * n).from_address(ptr)
    return torch.frombuffer(buf, dtype=dtype).view(*shp), gy, gz


def field_view(domain, name: str, rank: int = None):
    """interior view [nz][ny][nx] of one field of the current snapshot (no copy)."""
    t, gy, gz = snapshot_tensor(domain, rank)
    i = domain.model.fields.index(name)
    f = t[i]
    return f[gz:f.shape[0] - gz, gy:f.shape[1] - gy, :]


# ---------------------------------------------------------------------------------------------------------- DLPack
# Fields exported through the DLPack protocol (__dlpack__ / __dlpack_device__), so torch.from_dlpack(), cupy.from_dlpack()
# or jax.dlpack take a lattice field without a copy and without knowing this package.  The capsule is built with ctypes:
# a DLManagedTensor whose shape/strides describe the interior of one field inside the ghost-padded slab.
import ctypes as _C


class _DLDevice(_C.Structure):
    _fields_ = [("device_type", _C.c_int32), ("device_id", _C.c_int32)]


class _DLDataType(_C.Structure):
    _fields_ = [("code", _C.c_uint8), ("bits", _C.c_uint8), ("lanes", _C.c_uint16)]


class _DLTensor(_C.Structure):
    _fields_ = [("data", _C.c_void_p), ("device", _DLDevice), ("ndim", _C.c_int32), ("dtype", _DLDataType),
                ("shape", _C.POINTER(_C.c_int64)), ("strides", _C.POINTER(_C.c_int64)), ("byte_offset", _C.c_uint64)]


class _DLManagedTensor(_C.Structure):
    pass


_DLDeleter = _C.CFUNCTYPE(None, _C.POINTER(_DLManagedTensor))
_DLManagedTensor._fields_ = [("dl_tensor", _DLTensor), ("manager_ctx", _C.c_void_p), ("deleter", _DLDeleter)]
_live = {}          # address of a DLManagedTensor -> everything that must outlive it


@_DLDeleter
def _dl_deleter(p):
    _live.pop(_C.addressof(p.contents), None)


_capsule_destructor_t = _C.CFUNCTYPE(None, _C.c_void_p)


@_capsule_destructor_t
def _capsule_destructor(capsule):
    # a capsule that was never consumed still owns the tensor
    api = _C.pythonapi
    api.PyCapsule_IsValid.argtypes = [_C.c_void_p, _C.c_char_p]
    api.PyCapsule_IsValid.restype = _C.c_int
    if api.PyCapsule_IsValid(capsule, b"dltensor"):
        api.PyCapsule_GetPointer.argtypes = [_C.c_void_p, _C.c_char_p]
        api.PyCapsule_GetPointer.restype = _C.c_void_p
        _live.pop(api.PyCapsule_GetPointer(capsule, b"dltensor"), None)


class LatticeField:
    """One field of a rank's current snapshot as a DLPack producer: ``torch.from_dlpack(LatticeField(domain, "f[0]"))`` is a
    [nz][ny][nx] tensor over the lattice memory itself (writes land in the lattice).  The view stays valid while the
    domain lives and the snapshots do not swap (i.e. until the next iterate())."""

    def __init__(self, domain, name: str, rank: int = None):
        self.domain, self.name = domain, name
        self.rank = domain.local[0] if rank is None else rank
        ptr, shp, gy, gz, _ = domain.snapshot_info(self.rank, -1)
        self.esize = int(domain.model.d.storage_size)
        nf, NZ, NY, nx = (int(v) for v in shp)
        i = domain.model.fields.index(name)
        self.shape = (NZ - 2 * gz, NY - 2 * gy, nx)
        self.strides = (NY * nx, nx, 1)                           # in elements
        self.ptr = int(ptr) + self.esize * ((i * NZ + gz) * NY + gy) * nx
        self.on_device = domain.backend == 1
        self.device = domain.rank_device(self.rank) if hasattr(domain, "rank_device") else 0

    def __dlpack_device__(self):
        return (2, self.device) if self.on_device else (1, 0)     # kDLCUDA / kDLCPU

    def __dlpack__(self, stream=None, **kw):
        if self.on_device:
            self.domain.sync()                                    # the lattice stream is not the consumer's stream
        m = _DLManagedTensor()
        shape = (_C.c_int64 * 3)(*self.shape)
        strides = (_C.c_int64 * 3)(*self.strides)
        m.dl_tensor.data = self.ptr
        m.dl_tensor.device = _DLDevice(*self.__dlpack_device__())
        m.dl_tensor.ndim = 3
        m.dl_tensor.dtype = _DLDataType(2, 8 * self.esize, 1)     # kDLFloat
        m.dl_tensor.shape = shape
        m.dl_tensor.strides = strides
        m.dl_tensor.byte_offset = 0
        m.manager_ctx = None
        m.deleter = _dl_deleter
        addr = _C.addressof(m)
        _live[addr] = (m, shape, strides, self.domain)
        api = _C.pythonapi
        api.PyCapsule_New.argtypes = [_C.c_void_p, _C.c_char_p, _capsule_destructor_t]
        api.PyCapsule_New.restype = _C.py_object
        return api.PyCapsule_New(addr, b"dltensor", _capsule_destructor)
