"""Multi-process plumbing on CPU (gloo, world_size 2): halo exchange and Globals reduction through
torch.distributed hooks must reproduce the single-process result (BASELINE.json configs[0])."""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys, numpy as np
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
from tclb_b200.runtime import Model
from tclb_b200.parallel.dist import DistDomain
from helpers import channel_flags
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
m = Model("d2q9", "double")
shape = (64, 40, 1)
dd = DistDomain(m, shape, rank=rank, world=world, backend="cpu", direct=False)
d = dd.domain
fl = channel_flags(m, shape)
x0, y0, z0, nx, ny, nz = d.boxes[rank]
dd.set_local_flags(fl[z0:z0+nz, y0:y0+ny, x0:x0+nx])
d.set_setting("Viscosity", 0.05); d.set_setting("GravitationX", 1e-5)
d.init(); d.iterate(30, globals_on_last=True)
u = np.zeros((ny, nx, 3))
import ctypes as C
part = np.zeros((nz, ny, nx, 3)); 
from tclb_b200.runtime import _check
_check(d.lib.tclb_domain_get_quantity(d.h, rank, b"U", None, part.ctypes.data, 1.0))
np.save(os.path.join(%(out)r, "u_%%d.npy" %% rank), part)
np.save(os.path.join(%(out)r, "box_%%d.npy" %% rank), np.array(d.boxes[rank]))
dd.close()
'''


def test_two_process_gloo_matches_single(tmp_path, built):
    from tclb_b200 import build
    build.build_model_lib("d2q9", "double")
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import run_case
    ref = run_case("d2q9", "double", "cpu", (64, 40, 1), 30)
    script = tmp_path / "w.py"
    script.write_text(WORKER % dict(root=ROOT, out=str(tmp_path)))
    env = dict(os.environ, OMP_NUM_THREADS="2")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29611", str(script)], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    u = np.zeros_like(ref["U"])
    for rk in range(2):
        x0, y0, z0, nx, ny, nz = np.load(tmp_path / f"box_{rk}.npy")
        u[z0:z0 + nz, y0:y0 + ny, x0:x0 + nx] = np.load(tmp_path / f"u_{rk}.npy")
    assert np.array_equal(u, ref["U"])
