"""One process per GPU: torch.distributed does the plumbing (rendezvous, barriers, handle exchange),
the halo exchange and the Globals reduction themselves run INSIDE the stream-collide kernel through
CUDA-IPC mapped peer memory (NVLink/NVSwitch).  `direct=False` selects the baseline instead: the kernel
only updates its own subdomain and the ghost layers / Globals travel through NCCL (or gloo on CPU),
which is what the reference does with MPI (src/Lattice.cu.Rt:327-389, :1279-1292).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import numpy as np

from ..runtime import BACKEND_CUDA, Domain, Model, NativeError, _check, core

_EX_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_int)
_RED_FN = C.CFUNCTYPE(None, C.c_void_p, C.POINTER(C.c_double), C.c_int, C.c_int)


class DistDomain:
    def __init__(self, model: Model, shape, rank: int = 0, world: int = 1, device: int = 0, backend: str = "cuda",
                 direct: bool = True, init_pg: bool = True, domain: Optional[Domain] = None):
        """`domain`: wire an existing Domain (created with nranks=world, local_ranks=[rank]) instead of making one - how a
        case-file Solver joins a multi-process run (tclb_b200.run under torchrun)."""
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank, self.world, self.device, self.backend = rank, world, device, backend
        self.model = model
        self.direct = direct
        if world > 1 and init_pg and not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29500")
            kw = {}
            if backend == "cuda":
                kw["device_id"] = torch.device("cuda", device)
            dist.init_process_group("nccl" if backend == "cuda" else "gloo", rank=rank, world_size=world, **kw)
        self.domain = domain if domain is not None else Domain(model, shape, backend, nranks=world, local_ranks=[rank], devices=[device])
        self._cb = []
        if world > 1:
            if direct and backend == "cuda":
                self._wire_ipc()
                self._wire_hooks(exchange_too=False)     # handler collectives (Failcheck ...) still need the reduction
            else:
                self.domain.set_halo_mode(False)
                self._wire_hooks()

    # ---- fused path: map every rank's arena, let the kernels store into it ----
    def _wire_ipc(self):
        d, lib = self.domain, self.domain.lib
        handle = (C.c_ubyte * 64)()
        offs = (C.c_int64 * 4)()
        _check(lib.tclb_domain_export_ipc(d.h, self.rank, handle, offs))
        mine = (bytes(handle), list(offs), d.boxes[self.rank][4], d.boxes[self.rank][5])
        everyone = [None] * self.world
        self.dist.all_gather_object(everyone, mine)
        for p, (hb, po, pny, pnz) in enumerate(everyone):
            if p == self.rank:
                continue
            hbuf = (C.c_ubyte * 64).from_buffer_copy(hb)
            poffs = (C.c_int64 * 4)(*po)
            _check(lib.tclb_domain_import_peer(d.h, self.rank, p, hbuf, poffs, int(pny), int(pnz)))
        _check(lib.tclb_domain_finalize_wiring(d.h))
        self.barrier()

    # ---- baseline: NCCL / gloo ----
    def _wire_hooks(self, exchange_too: bool = True):
        torch, dist = self.torch, self.dist
        d = self.domain
        dev = torch.device("cuda", self.device) if self.backend == "cuda" else torch.device("cpu")
        nf = len(self.model.fields)
        gy, gz = self.model.gy, self.model.gz
        x0, y0, z0, nx, ny, nz = d.boxes[self.rank]
        # the snapshot holds STORAGE values (float16 for *_half builds), not the arithmetic type
        dtype = {2: torch.float16, 4: torch.float32, 8: torch.float64}[int(self.model.d.storage_size)]
        esize = int(self.model.d.storage_size)

        def snapshot_tensor(snap):
            ptr, shp, _, _, _ = d.snapshot_info(self.rank, snap)
            n = int(np.prod(shp))
            if self.backend == "cuda":
                from ..utils.interop import tensor_from_ptr
                return tensor_from_ptr(ptr, shp, dtype, self.device)
            buf = (C.c_char * (n * esize)).from_address(ptr)
            return torch.frombuffer(buf, dtype=dtype).view(*shp)

        divy, divz = d.divy, d.divz
        iy, iz = self.rank % divy, self.rank // divy

        def neighbour(dy, dz):
            return ((iy + dy) % divy) + divy * ((iz + dz) % divz)

        def exchange(_user, snap):
            if self.backend == "cuda":
                torch.cuda.current_stream(self.device).wait_stream(torch.cuda.ExternalStream(d.stream(self.rank), device=self.device))
                d.sync()
            t = snapshot_tensor(snap)             # [nf][NZ][NY][nx]
            ops, recvs = [], []
            for dz in (-1, 0, 1):
                for dy in (-1, 0, 1):
                    if (dy == 0 and dz == 0) or (dy != 0 and gy == 0) or (dz != 0 and gz == 0):
                        continue
                    nb = neighbour(dy, dz)
                    # rows I send towards (dy,dz); they land in the neighbour's ghost on its opposite side
                    ys = slice(gy + ny - gy, gy + ny) if dy > 0 else (slice(gy, 2 * gy) if dy < 0 else slice(gy, gy + ny))
                    zs = slice(gz + nz - gz, gz + nz) if dz > 0 else (slice(gz, 2 * gz) if dz < 0 else slice(gz, gz + nz))
                    send = t[:, zs, ys, :].contiguous()
                    # ghost region I receive from the neighbour at (-dy,-dz)... handled symmetrically by tag = direction index
                    yr = slice(0, gy) if dy > 0 else (slice(gy + ny, 2 * gy + ny) if dy < 0 else slice(gy, gy + ny))
                    zr = slice(0, gz) if dz > 0 else (slice(gz + nz, 2 * gz + nz) if dz < 0 else slice(gz, gz + nz))
                    src = neighbour(-dy, -dz)
                    recv = torch.empty_like(t[:, zr, yr, :].contiguous())
                    if nb == self.rank and src == self.rank:
                        t[:, zr, yr, :] = send
                        continue
                    tag = (dy + 1) + 3 * (dz + 1)
                    ops.append(dist.P2POp(dist.isend, send, nb, tag=tag))
                    ops.append(dist.P2POp(dist.irecv, recv, src, tag=tag))
                    recvs.append((recv, zr, yr))
            if ops:
                for w in dist.batch_isend_irecv(ops):
                    w.wait()
            for recv, zr, yr in recvs:
                t[:, zr, yr, :] = recv
            if self.backend == "cuda":
                torch.cuda.current_stream(self.device).synchronize()

        def reduce(_user, vals, n, op):
            a = np.ctypeslib.as_array(vals, shape=(n,))
            tt = torch.from_numpy(a.copy()).to(dev)
            dist.all_reduce(tt, op=dist.ReduceOp.SUM if op == 0 else dist.ReduceOp.MAX)
            a[:] = tt.cpu().numpy()

        ex, red = _EX_FN(exchange), _RED_FN(reduce)
        self._cb += [ex, red]
        _check(d.lib.tclb_domain_set_hooks(d.h, C.cast(ex, C.c_void_p) if exchange_too else None, C.cast(red, C.c_void_p), None))

    # ---- helpers ----
    def set_local_flags(self, flags_local: np.ndarray):
        fl = np.ascontiguousarray(flags_local, dtype=np.uint32)
        _check(self.domain.lib.tclb_domain_set_flags(self.domain.h, self.rank, fl.ctypes.data))

    def barrier(self):
        if self.world > 1:
            if self.backend == "cuda":
                self.dist.barrier(device_ids=[self.device])
            else:
                self.dist.barrier()

    def allreduce_max(self, v: float) -> float:
        if self.world == 1:
            return float(v)
        t = self.torch.tensor([float(v)], dtype=self.torch.float64,
                              device=self.torch.device("cuda", self.device) if self.backend == "cuda" else "cpu")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def e2e_bytes_per_step(self):
        lib = self.domain.lib
        lib.tclb_domain_io_bytes.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        # prime both paths once so the sizes are known
        zs = self.model.zonesettings
        self.domain.set_setting("Velocity" if "Velocity" in zs else ("VelocityX" if "VelocityX" in zs else zs[0]), 0.05)
        self.domain.iterate(1, globals_every=True)
        self.domain.globals()
        a, b = C.c_int64(), C.c_int64()
        _check(lib.tclb_domain_io_bytes(self.domain.h, self.rank, C.byref(a), C.byref(b)))
        return int(a.value), int(b.value)

    def close(self):
        self.domain.sync()
        self.barrier()
        self.domain.close()
        if self.world > 1 and self.dist.is_initialized():
            self.dist.destroy_process_group()
