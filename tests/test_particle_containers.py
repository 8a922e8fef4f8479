"""Particle containers (reference src/SolidAll.h, src/SolidGrid.h/.hpp; unit test tests/solid/main.cpp): random balls, the
uniform-grid container built next to the particle table must hand every node the same particles as the scan over all of
them - checked through the physics: fluid log and per-particle forces of the PSM model agree to rounding."""
import csv
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _case(n, shape, seed, iters, rmin=1.2, rmax=3.0):
    rng = np.random.default_rng(seed)
    nx, ny, nz = shape
    parts = "".join(f'<Particle x="{rng.uniform(0, nx):.4f}" y="{rng.uniform(0, ny):.4f}" z="{rng.uniform(0, nz):.4f}" r="{rng.uniform(rmin, rmax):.4f}" '
                    f'vx="{rng.uniform(-2e-3, 2e-3):.6f}" vy="{rng.uniform(-2e-3, 2e-3):.6f}" log="y"/>\n' for _ in range(n))
    return f"""<?xml version="1.0"?>
<CLBConfig version="2.0" permissive="true" output="output/">
  <Geometry nx="{nx}" ny="{ny}" nz="{nz}"><BGK><Box/></BGK></Geometry>
  <Model>
    <Param name="nu" value="0.03"/>
    <RemoteForceInterface integrator="SIMPLEPART">
      <SimplePart dt="1">
        {parts}
        <Periodic x="{nx}" y="{ny}" z="{nz}"/>
        <Log name="output/forces.csv" Iterations="1"/>
      </SimplePart>
    </RemoteForceInterface>
  </Model>
  <Log Iterations="2"/>
  <Solve Iterations="{iters}"/>
</CLBConfig>
"""


def _run(tclb_bin, wd, backend, container, xml, ranks=1):
    wd.mkdir()
    (wd / "p.xml").write_text(xml)
    env = dict(os.environ, TCLB_SOLID_CONTAINER=container)
    cmd = [tclb_bin, f"--{backend}", "--precision", "double"] + (["--ranks", str(ranks)] if ranks > 1 else []) + ["ref_d3q27_PSM_NEBB", "p.xml"]
    r = subprocess.run(cmd, cwd=wd, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return [list(csv.reader(open(wd / "output" / f))) for f in ("forces.csv", "p_Log_P00_00000000.csv")], r.stdout


def _compare(a, b, tol):
    for ta, tb in zip(a, b):
        assert ta[0] == tb[0] and len(ta) == len(tb) > 3
        for ra, rb in zip(ta[1:], tb[1:]):
            for name, x, y in zip(ta[0], ra, rb):
                if name == "Walltime":
                    continue
                x, y = float(x), float(y)
                assert abs(x - y) <= tol * max(abs(x), abs(y)) + 1e-13, (name, x, y)


@pytest.fixture
def psm_lib(built, reference_dir):
    from tclb_b200 import build
    build.build_from_dir(os.path.join(reference_dir, "models/particles/d3q27_PSM"), precision="double", options={"NEBB": True})


@pytest.mark.parametrize("ranks", [1, 2])
@pytest.mark.parametrize("container", ["grid", "tree"])
def test_containers_equal_scan_over_all_particles_cpu(psm_lib, tclb_bin, tmp_path, ranks, container):
    xml = _case(120, (32, 24, 16), 7, 6)
    (all_, _), (cont, out) = _run(tclb_bin, tmp_path / "all", "cpu", "all", xml, ranks), _run(tclb_bin, tmp_path / container, "cpu", container, xml, ranks)
    _compare(all_, cont, 1e-10)
    fx = [abs(float(r[8])) for r in cont[0][1:]]
    assert max(fx) > 1e-4           # the particles do exchange momentum with the fluid


def test_tree_with_very_different_radii_cpu(psm_lib, tclb_bin, tmp_path):
    """two large particles among many small ones: the case the tree exists for (the grid's cells grow with the largest)"""
    xml = _case(60, (32, 24, 16), 5, 6, rmin=0.8, rmax=1.2).replace("<Periodic", '<Particle x="10" y="12" z="8" r="6.5" log="y"/>\n<Particle x="24" y="10" z="7" r="5.0" vx="0.001" log="y"/>\n<Periodic', 1)
    (all_, _), (tree, _) = _run(tclb_bin, tmp_path / "all", "cpu", "all", xml), _run(tclb_bin, tmp_path / "tree", "cpu", "tree", xml)
    _compare(all_, tree, 1e-10)


@pytest.mark.gpu
def test_grid_container_gpu_matches_cpu_and_is_faster_than_the_scan(tclb_bin, tmp_path):
    if not os.path.exists(os.path.join(ROOT, "tclb_b200", "_lib", "libtclb_ref_d3q27_PSM_NEBB_double.so")):
        pytest.skip("ref_d3q27_PSM_NEBB was not built")
    xml = _case(150, (32, 24, 16), 11, 6)
    (cpu, _), (gpu, _) = _run(tclb_bin, tmp_path / "cpu", "cpu", "grid", xml), _run(tclb_bin, tmp_path / "gpu", "gpu", "grid", xml)
    _compare(cpu, gpu, 1e-8)
    (gtree, _) = _run(tclb_bin, tmp_path / "gpu_tree", "gpu", "tree", xml)
    _compare(cpu, gtree, 1e-8)
    # 10^4 particles on 128^3: the grid container against the scan over every particle
    import re
    import time
    big = _case(10000, (128, 128, 128), 3, 60, rmin=1.0, rmax=2.0).replace('<Log name="output/forces.csv" Iterations="1"/>', '<Log name="output/forces.csv" Iterations="20"/>').replace('<Log Iterations="2"/>', '<Log Iterations="20"/>')
    t = {}
    for container in ("grid", "tree", "all"):
        t0 = time.time()
        _run(tclb_bin, tmp_path / f"big_{container}", "gpu", container, big)
        t[container] = time.time() - t0
    print("10^4 particles, 128^3, 60 iterations: grid %.2f s, tree %.2f s, all %.2f s" % (t["grid"], t["tree"], t["all"]))
    assert t["all"] > 3 * t["grid"] and t["all"] > 2 * t["tree"], t
    # the same without process start-up and initialisation: a run four times as long, rate from the difference
    big4 = big.replace('<Solve Iterations="60"/>', '<Solve Iterations="240"/>')
    assert big4 != big
    rate = {}
    for container in ("grid", "tree", "all"):
        t0 = time.time()
        _run(tclb_bin, tmp_path / f"big4_{container}", "gpu", container, big4)
        rate[container] = 180.0 / max(time.time() - t0 - t[container], 1e-3)
    print("10^4 particles, 128^3: iterations per second  grid %.0f, tree %.0f, all %.0f  (grid/all = %.1fx, tree/all = %.1fx)"
          % (rate["grid"], rate["tree"], rate["all"], rate["grid"] / rate["all"], rate["tree"] / rate["all"]))
    assert rate["grid"] > 5 * rate["all"], rate
