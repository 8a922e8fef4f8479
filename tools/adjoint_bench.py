#!/usr/bin/env python
"""Throughput of the adjoint sweep next to the primal iteration (same lattice, same device).
Usage: adjoint_bench.py [model] [precision] [nx] [ny] [nz] [iterations]"""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from tclb_b200 import build  # noqa: E402
from tclb_b200.runtime import Adjoint, Domain, Model  # noqa: E402

model = sys.argv[1] if len(sys.argv) > 1 else "d2q9"
prec = sys.argv[2] if len(sys.argv) > 2 else "float"
nx, ny, nz = (int(v) for v in (sys.argv[3:6] if len(sys.argv) > 5 else (2048, 2048, 1)))
n = int(sys.argv[6]) if len(sys.argv) > 6 else 20
import os  # noqa: E402
if os.environ.get("AD_DIRS"):          # experiment: rebuild the adjoint companion with another chunk width first
    build.build_adjoint_lib(model, prec, dirs=int(os.environ["AD_DIRS"]))
import torch  # noqa: E402
backend = "cuda" if torch.cuda.is_available() else "cpu"
m = Model(model, prec)
d = Domain(m, (nx, ny, nz), backend, nranks=1)
fl = np.full((nz, ny, nx), m.nodetypes["MRT"][0], dtype=np.uint32)
fl[:, 0, :] = m.nodetypes["Wall"][0]
fl[:, -1, :] = m.nodetypes["Wall"][0]
d.set_flags(fl)
for k in m.zonesettings:
    if k.endswith("InObj"):
        d.set_setting(k, 1.0)
d.init()
d.iterate(10)
d.sync()
t0 = time.perf_counter(); d.iterate(n); d.sync(); tp = time.perf_counter() - t0
adj = Adjoint(d)
adj.record(n); adj.reverse(); adj.zero(); d.sync()          # warm-up: allocates the tape frames (recycled afterwards)
t0 = time.perf_counter(); adj.record(n); d.sync(); tr = time.perf_counter() - t0
t0 = time.perf_counter(); adj.reverse(); d.sync(); ta = time.perf_counter() - t0
nodes = nx * ny * nz * n
print(f"{model} {prec} {nx}x{ny}x{nz} on {backend}: primal {nodes / tp / 1e6:.0f} MLUPS, recording {nodes / tr / 1e6:.0f} MLUPS, "
      f"adjoint sweep {nodes / ta / 1e6:.0f} MLUPS (objective {adj.objective:.6g})")
