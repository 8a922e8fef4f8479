"""Design elements (reference src/Handlers/OptimalControl.cpp, OptimalControlSecond.cpp, Fourier.cpp, BSpline.cpp,
RepeatControl.cpp, InternalTopology.cpp, conFieldParameter.cpp, conExtrude.cpp) and the optimisers on top of them:
every design passes the finite-difference test of its gradient (<FDTest>, adjoint vs central differences), and every
built-in <Optimize> method raises the objective."""
import os
import re
import subprocess

import pytest

from tclb_b200 import build

HERE = os.path.dirname(os.path.abspath(__file__))

CONTROL_CASE = """<?xml version="1.0"?>
<CLBConfig version="2.0" output="output/">
  <Geometry nx="32" ny="16">
    <MRT><Box/></MRT>
    <WVelocity name="Inlet"><Inlet/></WVelocity>
    <EPressure name="Outlet"><Outlet/></EPressure>
    <Inlet nx="1" dx="3"><Box/></Inlet>
    <Outlet nx="1" dx="-3"><Box/></Outlet>
    <Wall mask="ALL"><Channel/></Wall>
  </Geometry>
  <Model>
    <Param name="VelocityX" value="0.01"/>
    <Param name="Viscosity" value="0.05"/>
  </Model>
  <Solve Iterations="24"/>
  <Control Iterations="8">
    <CSV file="ctl.csv" Time="t*1">
      <Param name="VelocityX" value="v*1" zone="Inlet"/>
    </CSV>
  </Control>
  {design}
  <Objective OutletFlux="1" PressureLoss="-0.5"/>
  <FDTest Epsilon="1e-7" {fd}>
    <Adjoint type="unsteady"><Solve Iterations="8"/></Adjoint>
  </FDTest>
  {tail}
</CLBConfig>
"""

DESIGNS = {
    "OptimalControl": ('<OptimalControl what="VelocityX-Inlet" lower="0" upper="0.05"/>', 8),
    "OptimalControlSecond": ('<OptimalControlSecond what="VelocityX-Inlet" lower="0" upper="0.05"/>', 4),
    "Fourier": ('<Fourier modes="3" lower="0" upper="0.05"><OptimalControl what="VelocityX-Inlet"/></Fourier>', 3),
    "BSpline": ('<BSpline nodes="5" order="2" lower="0" upper="0.05"><OptimalControl what="VelocityX-Inlet"/></BSpline>', 5),
    "BSplinePeriodic": ('<BSpline nodes="4" order="2" periodic="true" lower="0" upper="0.05"><OptimalControl what="VelocityX-Inlet"/></BSpline>', 4),
    "RepeatControl": ('<RepeatControl length="4" lower="0" upper="0.05"><OptimalControl what="VelocityX-Inlet"/></RepeatControl>', 4),
}


@pytest.fixture(scope="module")
def d2q9_libs(built):
    build.build_model_lib("d2q9", "double")
    build.build_adjoint_lib("d2q9", "double")


def _run(tclb_bin, tmp_path, model, xml, timeout=600, backend="cpu"):
    (tmp_path / "case.xml").write_text(xml)
    r = subprocess.run([tclb_bin, f"--{backend}", "--precision", "double", model, "case.xml"], cwd=tmp_path, capture_output=True, text=True, timeout=timeout)
    out = r.stdout + r.stderr
    assert r.returncode == 0, out[-4000:]
    return out


@pytest.mark.parametrize("name", sorted(DESIGNS))
def test_time_series_designs_pass_fdtest(d2q9_libs, tclb_bin, tmp_path, name):
    design, npar = DESIGNS[name]
    (tmp_path / "ctl.csv").write_text("t,v\n" + "".join(f"{t},{0.01 + 0.004 * ((3 * t) % 5)}\n" for t in range(9)))
    out = _run(tclb_bin, tmp_path, "d2q9", CONTROL_CASE.format(design=design, fd="", tail=""))
    assert f"Parameters in test: 0:{npar - 1} ({npar}) from {npar}" in out, out[-3000:]
    rows = re.findall(r"FDTest: parameter (\d+)  adjoint ([-+.eE0-9]+)  finite difference ([-+.eE0-9]+)", out)
    assert len(rows) == npar and {int(r[0]) for r in rows} == set(range(npar))
    assert any(abs(float(r[1])) > 1e-8 for r in rows)                     # the control does reach the objective
    # central differences of an objective of size ~1 with step 1e-7 carry ~1e-9 of rounding noise: parameters whose
    # gradient is (nearly) zero - the control of the last time level cannot reach the outlet any more - are compared absolutely
    for _, g, fd in rows:
        g, fd = float(g), float(fd)
        assert abs(g - fd) <= 1e-6 * max(abs(g), abs(fd)) + 5e-9, (name, g, fd)
    if name == "OptimalControl":
        trace = open(tmp_path / "output" / "case_OC_VelocityX_Inlet.dat").read()
        assert trace.startswith("GET,") and "GRAD," in trace and "SET," in trace


@pytest.mark.parametrize("method", ["MMA", "LBFGS", "GD", "NELDERMEAD"])
def test_optimize_methods_raise_the_objective_of_an_inlet_control(d2q9_libs, tclb_bin, tmp_path, method):
    (tmp_path / "ctl.csv").write_text("t,v\n" + "".join(f"{t},0.01\n" for t in range(9)))
    design = '<Fourier modes="3" lower="0" upper="0.03"><OptimalControl what="VelocityX-Inlet"/></Fourier>'
    tail = f'<Optimize method="{method}" MaxEvaluations="{9 if method != "NELDERMEAD" else 14}"><Adjoint type="unsteady"><Solve Iterations="8"/></Adjoint></Optimize>'
    out = _run(tclb_bin, tmp_path, "d2q9", CONTROL_CASE.format(design=design, fd='Samples="1"', tail=tail))
    ev = [float(v) for v in re.findall(r"Optimize: evaluation \d+ (?:best )?objective ([-+.eE0-9]+)", out)]
    best = float(re.search(r"Optimize: best objective ([-+.eE0-9]+)", out).group(1))
    assert f"Setting the algorithm to {method}" in out
    assert len(ev) >= 3 and best > ev[0] and best >= max(ev) - 1e-15, (ev, best)


TOPOLOGY_CASE = """<?xml version="1.0"?>
<CLBConfig version="2.0" output="output/" permissive="true">
  <Geometry nx="40" ny="20">
    <MRT><Box/></MRT>
    <WVelocity name="Inlet"><Inlet/></WVelocity>
    <EPressure name="Outlet"><Outlet/></EPressure>
    <Inlet nx="1" dx="3"><Box/></Inlet>
    <Outlet nx="1" dx="-3"><Box/></Outlet>
    <Wall mask="ALL"><Channel/></Wall>
    <DesignSpace><Box dx="14" nx="10" dy="6" ny="8"/></DesignSpace>
  </Geometry>
  <Model>
    <Param name="Velocity" value="0.01"/>
    <Param name="nu" value="0.05"/>
    <Param name="PorocityTheta" value="-3"/>
    <Param name="Porocity" value="0.4"/>
  </Model>
  <Solve Iterations="30"/>
  {design}
  <Objective Drag="-1.0" PressureLoss="0.2"/>
  <FDTest Samples="5" Epsilon="1e-6"><Adjoint type="unsteady"><Solve Iterations="6"/></Adjoint></FDTest>
  {tail}
</CLBConfig>
"""


@pytest.mark.parametrize("name,design,npar", [
    ("InternalTopology", "<InternalTopology/>", 80),
    ("FieldParameter", '<FieldParameter field="w" where="DesignSpace"/>', 80),
    ("Extrude", '<Extrude direction="x" theta="2.5" margin="1"><FieldParameter field="w" where="DesignSpace"/></Extrude>', 8),
])
def test_field_designs_pass_fdtest(reference_dir, tclb_bin, tmp_path, name, design, npar):
    build.build_from_dir(os.path.join(reference_dir, "models/optimization/d2q9_adj"), precision="double")
    build.build_adjoint_lib("ref_d2q9_adj", "double")
    tail = ""
    if name == "InternalTopology":
        tail = '<Optimize method="MMA" Material="less" MaxEvaluations="5"><Adjoint type="unsteady"><Solve Iterations="6"/></Adjoint></Optimize>'
    out = _run(tclb_bin, tmp_path, "ref_d2q9_adj", TOPOLOGY_CASE.format(design=design, tail=tail))
    assert f"from {npar}" in out, out[-3000:]
    worst = float(re.search(r"FDTest: largest relative difference ([-+.eE0-9]+)", out).group(1))
    assert worst < 1e-5, out[-3000:]
    if tail:
        ev = [float(v) for v in re.findall(r"Optimize: evaluation \d+ objective ([-+.eE0-9]+)", out)]
        best = float(re.search(r"Optimize: best objective ([-+.eE0-9]+)", out).group(1))
        assert len(ev) >= 3 and best >= ev[0]


@pytest.mark.gpu
def test_designs_pass_fdtest_on_the_gpu(tclb_bin, tmp_path):
    """the same finite-difference checks with the lattice, the tape and the adjoint kernels on the device: a time series
    design (OptimalControl of the inlet velocity) and a field design (InternalTopology of the reference's d2q9_adj)"""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    have = lambda n: os.path.exists(os.path.join(root, "tclb_b200", "_lib", n))
    if not (have("libtclb_d2q9_double_ad.so") and have("libtclb_ref_d2q9_adj_double_ad.so") and have("libtclb_ref_d2q9_adj_double.so")):
        pytest.skip("adjoint libraries were not built")
    a, b = tmp_path / "oc", tmp_path / "topo"
    a.mkdir(); b.mkdir()
    (a / "ctl.csv").write_text("t,v\n" + "".join(f"{t},{0.01 + 0.004 * ((3 * t) % 5)}\n" for t in range(9)))
    out = _run(tclb_bin, a, "d2q9", CONTROL_CASE.format(design=DESIGNS["OptimalControl"][0], fd="", tail=""), backend="gpu")
    rows = re.findall(r"FDTest: parameter (\d+)  adjoint ([-+.eE0-9]+)  finite difference ([-+.eE0-9]+)", out)
    assert len(rows) == 8 and any(abs(float(r[1])) > 1e-8 for r in rows)
    for _, g, fd in rows:
        g, fd = float(g), float(fd)
        assert abs(g - fd) <= 1e-6 * max(abs(g), abs(fd)) + 5e-9, (g, fd)
    out = _run(tclb_bin, b, "ref_d2q9_adj", TOPOLOGY_CASE.format(design="<InternalTopology/>", tail=""), backend="gpu")
    assert "from 80" in out
    worst = float(re.search(r"FDTest: largest relative difference ([-+.eE0-9]+)", out).group(1))
    assert worst < 1e-5, out[-3000:]
