"""Unmodified TCLB model directories through the R-subset interpreter + template engine (skipped without a mounted tree).

The strongest check: the reference's own d2q9 model files, lowered by this front-end and compiled by this generator,
reproduce the reference's golden Karman log."""
import csv
import os
import subprocess

import numpy as np
import pytest

from golden_karman import KARMAN_IT1000, KARMAN_XML

# models whose Dynamics.R + Dynamics.c[.Rt] must at least lower to C text
RENDER = ["PDE/diffusion2D", "flow/d2q9", "flow/d2q9_SRT", "flow/d2q9_les", "flow/d3q19", "flow/d3q27", "flow/d3q27_cumulant",
          "heat/d3q19_heat", "heat/d2q9q9_cm_cht", "multiphase/d3q27_pf_velocity", "multiphase/d2q9_kuper",
          "electrokinetic/d2q9_npe_guo", "electrokinetic/d2q9_poison_boltzmann", "shallowwater/sw"]


@pytest.mark.parametrize("path", RENDER)
def test_reference_model_lowers(reference_dir, path):
    from tclb_b200.gen import rfrontend
    m = rfrontend.load_model(os.path.join(reference_dir, "models", path))
    assert "CudaDeviceFunction" in m.code and len(m.fields) > 0 and len(m.stages) > 0
    assert "<?" not in m.code


def test_reference_d2q9_files_reproduce_karman_golden(reference_dir, tclb_bin, tmp_path):
    from tclb_b200 import build
    from tclb_b200.gen import rfrontend
    m = rfrontend.load_model(os.path.join(reference_dir, "models/flow/d2q9"), name="d2q9")
    build.build_custom(m, "ref_d2q9", "double")
    (tmp_path / "karman.xml").write_text(KARMAN_XML)
    r = subprocess.run([tclb_bin, "--cpu", "--precision", "double", "ref_d2q9", "karman.xml"], cwd=tmp_path, capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    rows = list(csv.reader(open(tmp_path / "output" / "karman_Log_P00_00000000.csv")))
    last = dict(zip(rows[0], rows[-1]))
    assert int(last["Iteration"]) == 1000
    for k, v in KARMAN_IT1000.items():
        assert abs(float(last[k]) - v) <= 1e-9 * abs(v), (k, last[k], v)


def test_native_library_matches_reference_tables(reference_dir):
    """the velocity sets and MRT matrices this framework generates equal the tables of src/lib/lattice.R / d3q19.R"""
    import numpy as np
    from tclb_b200.gen import rinterp, rlib
    I = rinterp.Interp(include_dirs=[os.path.join(reference_dir, "src")])
    I.run('source("lib/lattice.R")')          # no hooks installed here: this evaluates the reference's own file
    sets = rlib.velocity_sets()
    for name in ("d2q9", "d3q7", "d3q19", "d3q27"):
        assert np.array_equal(np.asarray(I.globals.get(name)).astype(int), sets[name]), name
    for name in ("d2q9", "d3q7", "d3q19"):
        assert rlib.known_matrix(sets[name]) is not None


@pytest.mark.parametrize("ranks", [1, 2])
def test_reference_pf_velocity_files_reproduce_drop_wall_golden(reference_dir, tclb_bin, tmp_path, ranks):
    """tests/external/d3q27_pf_velocity/drop_wall_120_{1,2}cpu.test: the two-lattice phase-field model (6 stages per
    Init, 4 per Iteration, dynamic stencil reads, 291 log columns) from the reference's own model files"""
    from tclb_b200 import build
    from tclb_b200.gen import rfrontend
    m = rfrontend.load_model(os.path.join(reference_dir, "models/multiphase/d3q27_pf_velocity"), name="d3q27_pf_velocity")
    build.build_custom(m, "ref_d3q27_pf_velocity", "double")
    case = os.path.join(reference_dir, "tests/external/d3q27_pf_velocity")
    (tmp_path / "drop_wall_120.xml").write_text(open(os.path.join(case, "drop_wall_120.xml")).read())
    r = subprocess.run([tclb_bin, "--cpu", "--precision", "double", "--ranks", str(ranks), "ref_d3q27_pf_velocity", "drop_wall_120.xml"],
                       cwd=tmp_path, capture_output=True, text=True, timeout=1200)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    d = subprocess.run(["python", os.path.join(root, "tools", "csvdiff"), "-a", str(tmp_path / "output" / "drop_wall_120_Log_P00_00000000.csv"),
                        "-b", os.path.join(case, "output", "drop_wall_120_Log_P00_00000000.csv"), "-x", "1e-8"], capture_output=True, text=True)
    assert d.returncode == 0, d.stdout


def test_reference_test_suite_files_pass_with_the_runner(reference_dir, tclb_bin):
    """tools/run_tests.py executes the reference's own `.test` scripts (karman, 2 ranks, restart, restart on 2 ranks,
    expected failure) against the native d2q9 model"""
    from tclb_b200 import build
    build.build_model_lib("d2q9", "double")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run(["python", os.path.join(root, "tools", "run_tests.py"), "d2q9", os.path.join(reference_dir, "tests/external/d2q9")],
                       capture_output=True, text=True, timeout=1800)
    assert r.returncode == 0 and "5 of 5 tests passed" in r.stdout, r.stdout[-3000:]


def test_reference_particle_suite_passes(reference_dir, tclb_bin):
    """tests/external/d3q27_PSM_NEBB: partially-saturated-cell particles (model lowered from the reference's files, NEBB
    option) coupled to the built-in SimplePart integrator: fluid log, particle force log and the VTK fields (with a
    periodic shift for the particle placed on the corner) against the goldens"""
    from tclb_b200 import build
    build.build_from_dir(os.path.join(reference_dir, "models/particles/d3q27_PSM"), precision="double", options={"NEBB": True})
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run(["python", os.path.join(root, "tools", "run_tests.py"), "d3q27_PSM_NEBB", os.path.join(reference_dir, "tests/external/d3q27_PSM_NEBB"),
                        "--library", "ref_d3q27_PSM_NEBB"], capture_output=True, text=True, timeout=1800)
    assert r.returncode == 0 and "3 of 3 tests passed" in r.stdout, r.stdout[-3000:]
    # the same suite with the launch lines taken literally: solver and `simplepart` as two programs (socket bridge)
    r = subprocess.run(["python", os.path.join(root, "tools", "run_tests.py"), "d3q27_PSM_NEBB", os.path.join(reference_dir, "tests/external/d3q27_PSM_NEBB"),
                        "--library", "ref_d3q27_PSM_NEBB", "--mpmd"], capture_output=True, text=True, timeout=1800)
    assert r.returncode == 0 and "3 of 3 tests passed" in r.stdout, r.stdout[-3000:]


def test_older_model_idioms_build_and_run(reference_dir, built):
    """model files written against earlier generations of the reference's node interface: the synthetic-turbulence time wave
    number read from `constContainer.ST.TimeWN` and the one-argument `SyntheticTurbulence(0)` (heat/experimental/
    d3q27_cumulant_heat), `for (ParticleIterator p(X,Y,Z); p; ++p)` (flow/d2q9_par)"""
    import numpy as np
    from tclb_b200 import build
    from tclb_b200.runtime import Domain, Model
    build.build_from_dir(os.path.join(reference_dir, "models/heat/experimental/d3q27_cumulant_heat"), precision="double")
    m = Model("ref_d3q27_cumulant_heat", "double")
    assert "WVelocityTurbulent" in m.nodetypes and "T" in m.quantities
    d = Domain(m, (16, 12, 8), "cpu")
    fl = np.full((8, 12, 16), m.flag("MRT"), dtype=np.uint32)
    fl[:, 0, :] = m.flag("Wall"); fl[:, -1, :] = m.flag("Wall")
    fl[:, 1:-1, 0] = m.flag("MRT", "WVelocityTurbulent")
    d.set_flags(fl)
    for k, v in (("nu", 0.05), ("ForceX", 1e-5), ("Velocity", 0.01)):
        if k in m.settings or k in m.zonesettings:
            d.set_setting(k, v)
    d.init(); d.iterate(20)
    u = d.get_quantity("U")
    assert np.isfinite(u).all() and np.abs(u).max() > 0
    d.close()
    src = build.build_from_dir(os.path.join(reference_dir, "models/flow/d2q9_par"), precision="double")
    assert os.path.exists(str(src))


def test_wide_flag_model_builds_and_runs_its_example(reference_dir, built, tclb_bin, tmp_path):
    """heat/d2q9q9_cm_cht needs a 32-bit flag word (8 node-type groups + zones): the library builds with nvcc (the catch-all
    group's capacity does not fit an int) and the reference's lid-driven cavity example runs"""
    import re
    from tclb_b200 import build
    from tclb_b200.runtime import Model
    build.build_from_dir(os.path.join(reference_dir, "models/heat/d2q9q9_cm_cht"), precision="double")
    m = Model("ref_d2q9q9_cm_cht", "double")
    assert m.flag_dtype.__name__ == "uint32"
    xml = open(os.path.join(reference_dir, "example/heat/d2q9q9_lid_driven_cavity.xml")).read()
    (tmp_path / "c.xml").write_text(re.sub(r'Iterations="\d+"', 'Iterations="4"', xml))
    r = subprocess.run([tclb_bin, "--cpu", "--precision", "double", "ref_d2q9q9_cm_cht", "c.xml"], cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


def test_cm_cht_cavity_keeps_a_uniform_temperature(reference_dir, built, tmp_path, monkeypatch):
    """heat/d2q9q9_cm_cht on its own lid-driven cavity example (example/heat/d2q9q9_lid_driven_cavity.xml, smaller box): with
    the lid at rest the uniform state is stationary to rounding; with the lid moving a flow develops, mass is conserved and the
    temperature (enthalpy lattice / density / cp) stays at its uniform initial value away from the lid corners"""
    import re
    from tclb_b200 import build
    from tclb_b200.solver import Solver
    build.build_from_dir(os.path.join(reference_dir, "models/heat/d2q9q9_cm_cht"), precision="double")
    xml = open(os.path.join(reference_dir, "example/heat/d2q9q9_lid_driven_cavity.xml")).read()
    xml = xml[xml.index("<?xml"):]
    xml = xml.replace('nx="256" ny="256"', 'nx="48" ny="48"').replace('<Box dx="37" dy="107" nx="50" ny="50"/>', '<Box dx="10" dy="20" nx="8" ny="8"/>')
    xml = re.sub(r'<Solve Iterations="3">.*?</Solve>', '', xml, flags=re.S)
    xml = re.sub(r'<Failcheck[^>]*/>', '', xml)
    xml = re.sub(r'<Solve Iterations="50000">.*?</Solve>', '<Solve Iterations="600"/>', xml, flags=re.S)
    assert 'Iterations="600"' in xml
    monkeypatch.chdir(tmp_path)

    def run(lid):
        s = Solver("ref_d2q9q9_cm_cht", "double", "cpu")
        s.load_string(xml.replace('name="VelocityX" value="0.0001"', f'name="VelocityX" value="{lid}"'), "cht").run()
        T, U, R = (np.asarray(s.Quantities[k]) for k in ("T", "U", "Rho"))
        s.close()
        return T[0], U[0], R[0]

    T, U, R = run(0)
    assert np.abs(T - 10).max() < 1e-12 and np.abs(U).max() < 1e-14 and np.abs(R - 1).max() < 1e-14
    T, U, R = run(1e-4)
    assert 5e-5 < np.abs(U).max() < 2e-4                         # the lid drives the cavity
    assert abs(R.sum() - R.size) < 1e-4                          # mass
    assert np.abs(T[4:-4, 4:-4] - 10).max() < 1e-2               # uniform temperature away from the walls
    dev = np.abs(T - 10)
    assert np.unravel_index(dev.argmax(), dev.shape)[0] >= T.shape[0] - 2    # what deviation there is sits at the lid


def _write_cylinder_stl(path, r, z0, z1, n=96):
    """binary STL of a closed cylinder around the z axis.  A ray that passes through a vertex (within float rounding) is a hit
    or a miss by rounding, in the reference's cut search as in this one (src/Geometry.cpp.Rt:479-503), so the polygon is turned
    away from the lattice directions."""
    import math
    import struct
    tris = []
    for i in range(n):
        a0, a1 = 2 * math.pi * (i + 0.37) / n, 2 * math.pi * (i + 1.37) / n      # no vertex on a lattice line or diagonal through the axis
        p0, p1 = (r * math.cos(a0), r * math.sin(a0)), (r * math.cos(a1), r * math.sin(a1))
        tris += [((p0[0], p0[1], z0), (p1[0], p1[1], z0), (p1[0], p1[1], z1)), ((p0[0], p0[1], z0), (p1[0], p1[1], z1), (p0[0], p0[1], z1)),
                 ((0, 0, z0), (p1[0], p1[1], z0), (p0[0], p0[1], z0)), ((0, 0, z1), (p0[0], p0[1], z1), (p1[0], p1[1], z1))]
    with open(path, "wb") as f:
        f.write(b"cylinder".ljust(80, b" "))
        f.write(struct.pack("<I", len(tris)))
        for t in tris:
            f.write(struct.pack("<3f", 0, 0, 0))
            for p in t:
                f.write(struct.pack("<3f", *p))
            f.write(struct.pack("<H", 0))


def _hot_karman(reference_dir, tmp_path, monkeypatch, backend, iterations=800):
    import re
    from tclb_b200 import build
    from tclb_b200.solver import Solver
    build.build_from_dir(os.path.join(reference_dir, "models/heat/d3q27q7_cm_cht"), precision="double", options={"OutFlowNeumann": True, "AVG": True, "IBB": True})
    os.makedirs(tmp_path / "example" / "heat")
    _write_cylinder_stl(tmp_path / "example" / "heat" / "cylinder_XYplane_D10_Z9.stl", 5.0, -1.0, 9.0)
    xml = open(os.path.join(reference_dir, "example/heat/d3q27q7_Hot_Karman_Re100_Pr5.xml")).read()
    xml = xml[xml.index("<?xml"):]
    xml = re.sub(r'<Solve Iterations="3">.*?</Solve>', '', xml, flags=re.S)
    xml = re.sub(r'<Solve Iterations="30000">.*?</Solve>', '<Solve Iterations="%d"/>' % iterations, xml, flags=re.S)
    xml = xml.replace('Failcheck Iterations="5000"', 'Failcheck Iterations="100"')
    monkeypatch.chdir(tmp_path)
    s = Solver("ref_d3q27q7_cm_cht_OutFlowNeumann_AVG_IBB", "double", backend)
    s.load_string(xml, "hk").run()
    T, U, R = (np.asarray(s.Quantities[k]) for k in ("T", "U", "Rho"))
    s.close()
    return T, U, R


def test_cm_cht_3d_heated_cylinder_with_interpolated_bounce_back(reference_dir, built, tmp_path, monkeypatch):
    """heat/d3q27q7_cm_cht built with the options its example names (OutFlowNeumann, AVG, IBB; the node code declares `cut_t`
    variables) on example/heat/d3q27q7_Hot_Karman_Re100_Pr5.xml: wall cuts from an STL surface (the example's STL is not in the
    tree, a cylinder of the same size is written here), interpolated bounce-back for both lattices, Dirichlet temperature on the
    cylinder.  The temperature obeys the maximum principle between inflow (10) and cylinder (11) and the wake warms up."""
    T, U, R = _hot_karman(reference_dir, tmp_path, monkeypatch, "cpu")
    assert T.shape == (3, 100, 512) and np.isfinite(T).all()
    assert 9.9 < T.min() and T.max() < 11.1 and T.max() > 10.9           # maximum principle (small overshoot of the scheme)
    assert abs(R.mean() - 1) < 5e-3 and 0.03 < U[..., 0].max() < 0.06     # inflow 0.025 accelerates around the cylinder
    assert T[1, 45:56, 152:170].mean() > 10.05                           # the wake right behind the cylinder (x 120..150) is heated
    assert abs(T[1, 5:15, 300:500].mean() - 10) < 5e-3                   # far away nothing has arrived yet
    assert np.abs(T[0] - T[2]).max() < 1e-9                              # the case is uniform along z


HEATED_CYLINDER = """<?xml version="1.0"?>
<CLBConfig version="2.0" output="output/">
  <Geometry nx="192" ny="64" nz="3">
    <CM><Box/></CM>
    <WVelocity name="Inlet"><Box nx="1"/></WVelocity>
    <ENeumann name="Outlet"><Box dx="-1"/></ENeumann>
    <HydroIBB><STL file="cyl.stl" scale="2.0" x="60.3" y="32.2" z="0" side="surface"/></HydroIBB>
    <HeaterDirichletTemperatureIABB name="Hot"><STL file="cyl.stl" scale="2.0" x="60.3" y="32.2" z="0" side="surface"/></HeaterDirichletTemperatureIABB>
  </Geometry>
  <Model>
    <Param name="InitTemperature" value="10"/>
    <Param name="InitTemperature" value="10" zone="Inlet"/>
    <Param name="InitTemperature" value="11" zone="Hot"/>
    <Param name="nu" value="0.02"/><Param name="conductivity" value="0.01"/>
    <Param name="nu_buffer" value="0.166666"/><Param name="conductivity_buffer" value="0.166666"/>
    <Param name="h_stability_enhancement" value="1."/><Param name="cp" value="1."/><Param name="material_density" value="1."/>
    <Param name="VelocityX" value="0.03"/>
  </Model>
  <Solve Iterations="300"/>
</CLBConfig>
"""


@pytest.mark.gpu
def test_cm_cht_3d_heated_cylinder_gpu_matches_cpu(built, tmp_path, monkeypatch):
    """d3q27q7_cm_cht (OutFlowNeumann, AVG, IBB) on the GPU - wall cuts in device memory, interpolated bounce-back of both
    lattices inside the fused stage kernel - against the CPU oracle of the same generated code.  The library is the one the CPU
    test builds from the reference's directory; the case is this repository's own, so the test also runs where the reference
    tree is not mounted."""
    from tclb_b200.runtime import LIBDIR
    from tclb_b200.solver import Solver
    name = "ref_d3q27q7_cm_cht_OutFlowNeumann_AVG_IBB"
    if not (LIBDIR / f"libtclb_{name}_double.so").exists():
        pytest.skip("library not built (tests/test_reference_frontend.py builds it where the reference tree is mounted)")
    _write_cylinder_stl(tmp_path / "cyl.stl", 5.0, -1.0, 9.0)
    monkeypatch.chdir(tmp_path)
    out = {}
    for backend in ("cpu", "cuda"):
        s = Solver(name, "double", backend)
        s.load_string(HEATED_CYLINDER, "hc_" + backend).run()
        out[backend] = [np.asarray(s.Quantities[k]) for k in ("T", "U", "Rho")]
        s.close()
    (Tc, Uc, Rc), (Tg, Ug, Rg) = out["cpu"], out["cuda"]
    assert Tc.max() > 10.9 and Tc.min() > 9.9 and Tc[1, 28:37, 75:90].mean() > 10.01
    assert np.abs(Tg - Tc).max() < 1e-9 and np.abs(Ug - Uc).max() < 1e-10 and np.abs(Rg - Rc).max() < 1e-10
