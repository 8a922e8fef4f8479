"""External particle integrators over the local-socket bridge (csrc/host/rfi_bridge.h; reference: RemoteForceInterface +
MPMD, src/RemoteForceInterface.hpp:596-735, side programs src/simplepart.cpp and src/empty.cpp): the integrator running as
ANOTHER PROGRAM (started by the solver, or listening on a socket beforehand) must give the logs of the in-process one."""
import csv
import os
import subprocess
import time

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INTEG = os.path.join(ROOT, "bin", "tclb-integrator")


def _simplepart(n, shape, seed):
    rng = np.random.default_rng(seed)
    nx, ny, nz = shape
    parts = "".join(f'<Particle x="{rng.uniform(0, nx):.4f}" y="{rng.uniform(0, ny):.4f}" z="{rng.uniform(0, nz):.4f}" r="{rng.uniform(1.5, 3.0):.4f}" '
                    f'vx="{rng.uniform(-2e-3, 2e-3):.6f}" vy="{rng.uniform(-2e-3, 2e-3):.6f}" log="y"/>\n' for _ in range(n))
    return f"""<SimplePart dt="1">
        {parts}
        <Periodic x="{nx}" y="{ny}" z="{nz}"/>
        <Log name="output/forces.csv" Iterations="1"/>
      </SimplePart>"""


def _case(shape, rfi, iters=6):
    nx, ny, nz = shape
    return f"""<?xml version="1.0"?>
<CLBConfig version="2.0" permissive="true" output="output/">
  <Geometry nx="{nx}" ny="{ny}" nz="{nz}"><BGK><Box/></BGK></Geometry>
  <Model>
    <Param name="nu" value="0.03"/>
    {rfi}
  </Model>
  <Log Iterations="2"/>
  <Solve Iterations="{iters}"/>
</CLBConfig>
"""


def _run(tclb_bin, wd, xml, ranks=1):
    (wd / "p.xml").write_text(xml)
    cmd = [tclb_bin, "--cpu", "--precision", "double"] + (["--ranks", str(ranks)] if ranks > 1 else []) + ["ref_d3q27_PSM_NEBB", "p.xml"]
    r = subprocess.run(cmd, cwd=wd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return r.stdout


def _tables(wd, names=("forces.csv", "p_Log_P00_00000000.csv")):
    return [list(csv.reader(open(wd / "output" / f))) for f in names]


def _same(a, b, tol=1e-10):
    """equal up to the order of the force atomics (last digits); the wall-clock column is skipped"""
    for ta, tb in zip(a, b):
        assert ta[0] == tb[0] and len(ta) == len(tb) > 3
        for ra, rb in zip(ta[1:], tb[1:]):
            for name, x, y in zip(ta[0], ra, rb):
                if name.strip() == "Walltime":
                    continue
                x, y = float(x), float(y)
                assert abs(x - y) <= tol * max(abs(x), abs(y)) + 1e-13, (name, x, y)


@pytest.fixture
def psm_lib(built, reference_dir):
    from tclb_b200 import build
    build.build_from_dir(os.path.join(reference_dir, "models/particles/d3q27_PSM"), precision="double", options={"NEBB": True})


@pytest.mark.parametrize("ranks", [1, 2])
def test_spawned_simplepart_equals_in_process(psm_lib, tclb_bin, tmp_path, ranks):
    shape = (32, 24, 16)
    sp = _simplepart(40, shape, 3)
    a, b = tmp_path / "inproc", tmp_path / "spawn"
    a.mkdir(); b.mkdir()
    _run(tclb_bin, a, _case(shape, f'<RemoteForceInterface integrator="SIMPLEPART">{sp}</RemoteForceInterface>'), ranks)
    (b / "sp.xml").write_text('<?xml version="1.0"?>\n' + sp)
    out = _run(tclb_bin, b, _case(shape, f'<RemoteForceInterface integrator="simplepart" spawn="{INTEG} simplepart sp.xml"/>'), ranks)
    assert "Connecting RFI to simplepart (starting:" in out
    ta, tb = _tables(a), _tables(b)
    _same(ta, tb)                        # same integrator code on the other side of a socket
    assert len(ta[0]) > 6 and max(abs(float(r[8])) for r in ta[0][1:]) > 1e-5


def test_listening_integrator_and_empty(psm_lib, tclb_bin, tmp_path):
    shape = (24, 16, 12)
    sock = str(tmp_path / "rfi.sock")
    # an integrator that was started first and listens: `empty` takes part in every exchange with no particles
    wd = tmp_path / "empty"
    wd.mkdir()
    srv = subprocess.Popen([INTEG, "empty", "--listen", sock], cwd=wd, stdout=subprocess.PIPE, text=True)
    try:
        _run(tclb_bin, wd, _case(shape, f'<RemoteForceInterface integrator="empty" connect="{sock}"/>'))
        out, _ = srv.communicate(timeout=60)
    finally:
        if srv.poll() is None:
            srv.kill()
    assert srv.returncode == 0 and "served 7 exchanges" in out   # the initialisation action + 6 iterations, out
    # no particles = the same fluid as an in-process integrator without particles
    ref = tmp_path / "none"
    ref.mkdir()
    _run(tclb_bin, ref, _case(shape, '<RemoteForceInterface integrator="SIMPLEPART"><SimplePart dt="1"><Periodic x="24" y="16" z="12"/></SimplePart></RemoteForceInterface>'))
    _same(_tables(wd, ("p_Log_P00_00000000.csv",)), _tables(ref, ("p_Log_P00_00000000.csv",)))


def test_missing_integrator_is_a_clear_error(psm_lib, tclb_bin, tmp_path):
    (tmp_path / "p.xml").write_text(_case((16, 8, 8), '<RemoteForceInterface integrator="lammps" spawn="/nonexistent/lmp in.file"/>'))
    r = subprocess.run([tclb_bin, "--cpu", "--precision", "double", "ref_d3q27_PSM_NEBB", "p.xml"], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode != 0 and "did not answer the handshake" in (r.stdout + r.stderr)
