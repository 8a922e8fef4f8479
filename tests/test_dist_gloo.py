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
d.init()
if %(perturb)d:
    # fields written from the host after initialisation: the neighbours' ghost rows must follow (exchange hook)
    gx, gy_ = np.meshgrid(np.arange(shape[0]), np.arange(shape[1]))
    for k, f in enumerate(m.fields[:3]):
        base = np.zeros((1, shape[1], shape[0])); base[0, y0:y0+ny, x0:x0+nx] = d.get_field(f)[0, y0:y0+ny, x0:x0+nx]
        d.set_field(f, base * (1.0 + 1e-3 * np.sin(0.3 * gx + 0.2 * gy_ + k))[None])
d.iterate(30, globals_on_last=True)
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
    script.write_text(WORKER % dict(root=ROOT, out=str(tmp_path), perturb=0))
    env = dict(os.environ, OMP_NUM_THREADS="2")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29611", str(script)], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    u = np.zeros_like(ref["U"])
    for rk in range(2):
        x0, y0, z0, nx, ny, nz = np.load(tmp_path / f"box_{rk}.npy")
        u[z0:z0 + nz, y0:y0 + ny, x0:x0 + nx] = np.load(tmp_path / f"u_{rk}.npy")
    assert np.array_equal(u, ref["U"])


def test_two_process_set_field_refreshes_neighbour_ghosts(tmp_path, built):
    from tclb_b200 import build
    from tclb_b200.runtime import Domain, Model
    build.build_model_lib("d2q9", "double")
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import setup_case
    m = Model("d2q9", "double")
    shape = (64, 40, 1)
    d = Domain(m, shape, "cpu")
    setup_case(d, m, shape)
    gx, gy_ = np.meshgrid(np.arange(shape[0]), np.arange(shape[1]))
    for k, f in enumerate(m.fields[:3]):
        d.set_field(f, d.get_field(f) * (1.0 + 1e-3 * np.sin(0.3 * gx + 0.2 * gy_ + k))[None])
    d.iterate(30, globals_on_last=True)
    ref = d.get_quantity("U")
    d.close()
    script = tmp_path / "w.py"
    script.write_text(WORKER % dict(root=ROOT, out=str(tmp_path), perturb=1))
    env = dict(os.environ, OMP_NUM_THREADS="2")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29613", str(script)], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    u = np.zeros_like(ref)
    for rk in range(2):
        x0, y0, z0, nx, ny, nz = np.load(tmp_path / f"box_{rk}.npy")
        u[z0:z0 + nz, y0:y0 + ny, x0:x0 + nx] = np.load(tmp_path / f"u_{rk}.npy")
    assert np.array_equal(u, ref)


def test_case_file_on_two_processes_matches_golden_karman(tmp_path, built):
    """The reference's karman case run as a CASE FILE by two processes (torch.distributed/gloo, one subdomain each,
    `python -m tclb_b200.run` under torchrun): rank 0 writes the Log (golden values of tests/external/d2q9/karman_p2.test)
    and the .pvti index, every rank its .vti piece; a checkpoint written on the way restarts to the same result."""
    import csv
    from tclb_b200 import build
    build.build_model_lib("d2q9", "double")
    from golden_karman import KARMAN_IT1000, KARMAN_XML
    xml = KARMAN_XML.replace("<Solve", '<VTK Iterations="1000"/>\n<SaveCheckpoint Iterations="500" keep="all"/>\n<Solve', 1)
    (tmp_path / "karman.xml").write_text(xml)
    env = dict(os.environ, OMP_NUM_THREADS="2", PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))

    def launch(case, port):
        r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                            "--master-port", str(port), "-m", "tclb_b200.run", "--cpu", "--precision", "double", "d2q9", case],
                           cwd=tmp_path, capture_output=True, text=True, timeout=1200, env=env)
        assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]

    launch("karman.xml", 29621)
    out = tmp_path / "output"
    rows = list(csv.reader(open(out / "karman_Log_P00_00000000.csv")))
    last = dict(zip(rows[0], rows[-1]))
    for k, v in KARMAN_IT1000.items():
        assert abs(float(last[k]) - v) <= 1e-9 * abs(v), (k, last[k], v)
    names = sorted(os.listdir(out))
    assert "karman_VTK_P00_00001000.pvti" in names and "karman_VTK_P00_00001000.vti" in names and "karman_VTK_P01_00001000.vti" in names
    assert b"karman_VTK_P01_00001000.vti" in open(out / "karman_VTK_P00_00001000.pvti", "rb").read()
    # restart from the checkpoint of iteration 500 (per-rank .pri files + restart XML written by rank 0)
    assert "karman_checkpoint_00000500_0.pri" in names and "karman_checkpoint_00000500_1.pri" in names and "karman_restart_00000500.xml" in names
    launch(os.path.join("output", "karman_restart_00000500.xml"), 29622)
    logs = [f for f in os.listdir(out) if f.startswith("karman_restart_00000500_Log")]
    assert logs, sorted(os.listdir(out))
    # the iteration counter is not part of a checkpoint (like the reference): 500 iterations after the restart = 1000 in all
    rows2 = list(csv.reader(open(out / logs[0])))
    at500 = [dict(zip(rows2[0], r)) for r in rows2[1:] if int(float(r[0])) == 500][0]
    for k, v in KARMAN_IT1000.items():
        assert abs(float(at500[k]) - v) <= 1e-9 * abs(v), (k, at500[k], v)
