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
        dev = 0
        return tensor_from_ptr(ptr, shp, dtype, torch.cuda.current_device()), gy, gz
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
