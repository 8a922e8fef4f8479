"""<RunPython> coupling (reference src/Handlers/cbRunR.cpp; tests/external/d2q9/pythonGeom.xml, d2q9_bc/runexternal.xml):
the code of the element runs in-process with a `Solver` object that reads and writes the live lattice."""
import numpy as np
import pytest

CASE = """<?xml version="1.0"?>
<CLBConfig version="2.0" output="output/">
  <Geometry nx="64" ny="32">
    <MRT><Box/></MRT>
    <WVelocity name="Inlet"><Inlet/></WVelocity>
    <EPressure name="Outlet"><Outlet/></EPressure>
    <Wall mask="ALL"><Channel/></Wall>
  </Geometry>
  <RunPython><![CDATA[
    x = (Solver.Geometry.X - 0.5) / 64
    y = (Solver.Geometry.Y - 0.5) / 16
    sel = y < np.sin(np.pi * x * 2) / 4 + 0.5
    Solver.Geometry["BOUNDARY"][sel] = "Wall"
    Solver.Geometry["COLLISION"][sel] = "None"
    painted = int(sel.sum())
    Solver.Settings["VelocityX", "Inlet"] = 0.02
  ]]></RunPython>
  <Model>
    <Param name="Viscosity" value="0.05"/>
  </Model>
  <RunPython Iterations="20">
    history.append((Solver.Info.Iteration, float(np.abs(Solver.Quantities["U"]).max())))
  </RunPython>
  <Solve Iterations="60"/>
  <RunPython>
    f3 = Solver.Fields["f[3]"]
    Solver.Fields["f[3]"] = f3 * 0 + 0.125
    after = Solver.Fields["f[3]"]
  </RunPython>
</CLBConfig>
"""


def test_runpython_edits_geometry_settings_and_fields(built, tmp_path, monkeypatch):
    from tclb_b200 import build
    from tclb_b200.solver import Solver
    build.build_model_lib("d2q9", "double")
    monkeypatch.chdir(tmp_path)
    s = Solver("d2q9", "double", "cpu")
    s.namespace["history"] = []
    s.load_string(CASE, "coupled").run()
    ns = s.namespace
    m = s.model
    fl = s.domain.get_flags()
    wall = (fl & m.nodetypes["Wall"][1]) == m.nodetypes["Wall"][0]
    assert ns["painted"] > 100 and wall.sum() >= ns["painted"]
    assert [it for it, _ in ns["history"]] == [20, 40, 60]
    assert 0 < ns["history"][-1][1] < 0.1                      # the inlet velocity set from Python drives a flow
    assert np.allclose(ns["after"], 0.125)
    s.close()


def test_runpython_without_python_host_fails_cleanly(tclb_bin, tmp_path):
    import subprocess
    (tmp_path / "c.xml").write_text(CASE)
    r = subprocess.run([tclb_bin, "--cpu", "d2q9", "c.xml"], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode != 0 and "tclb_b200.run" in (r.stdout + r.stderr)


R_CASE = """<?xml version="1.0"?>
<CLBConfig version="2.0" output="output/">
  <Geometry nx="32" ny="16">
    <MRT><Box/></MRT>
    <Wall mask="ALL"><Channel/></Wall>
  </Geometry>
  <Model>
    <Param name="Viscosity" value="0.05"/>
  </Model>
  <RunR><![CDATA[
    x = (Solver$Geometry$X - 0.5) / 32
    y = (Solver$Geometry$Y - 0.5) / 16
    sel = (x > 0.4) & (x < 0.6) & (y < 0.5)
    Solver$Geometry$BOUNDARY[sel] = "Wall"
    Solver$Geometry$COLLISION[sel] = "None"
  ]]></RunR>
  <Solve Iterations="10"/>
  <RunR>
    nu0 = Solver$Settings$Viscosity
    Solver$Settings$Viscosity = nu0 * 2
    f3 = Solver$Fields$f3
    x = (Solver$Geometry$X - 0.5) / 32
    Solver$Fields$f3[] = 0.1 + 0.01 * sin(2 * pi * x)
    Solver$Actions$Iterate(3)
    u = Solver$Quantities$U
    write.csv(data.frame(it = Solver$Iteration, umax = max(abs(u)), n = length(f3)), "output/r_probe.csv", row.names = FALSE)
  </RunR>
</CLBConfig>
"""


def test_runr_blocks_run_in_the_r_subset_interpreter(built, tmp_path, monkeypatch):
    """<RunR> (reference src/Handlers/cbRunR.cpp) without R: settings, fields, quantities, geometry coordinates through
    a `Solver` list; what the block changes is written back to the lattice"""
    from tclb_b200 import build
    from tclb_b200.solver import Solver
    build.build_model_lib("d2q9", "double")
    monkeypatch.chdir(tmp_path)
    s = Solver("d2q9", "double", "cpu")
    case = R_CASE
    fname = s.model.fields[3]
    case = case.replace("Solver$Fields$f3", "Solver$Fields$`%s`" % fname if not fname.isidentifier() else "Solver$Fields$" + fname)
    s.load_string(case, "rcase").run()
    assert abs(s.domain.get_setting("Viscosity") - 0.1) < 1e-15
    fl = s.domain.get_flags()
    wall = (fl & s.model.nodetypes["Wall"][1]) == s.model.nodetypes["Wall"][0]
    xs, ys = (np.arange(32) + 0.0) / 32, (np.arange(16) + 0.0) / 16
    painted = ((xs > 0.4) & (xs < 0.6))[None, :] & (ys < 0.5)[:, None]
    assert wall[0][painted].all() and not wall[0][1:-1][:, :8].any()              # the block painted by the R code, nothing else inside
    assert ((fl[0][painted] & s.model.nodegroups["COLLISION"][0]) == 0).all()
    f3 = s.domain.get_field(fname)
    x = (np.arange(32) + 0.5 - 0.5) / 32
    assert not np.allclose(f3[0], 0.1 + 0.01 * np.sin(2 * np.pi * x)[None, :], rtol=0, atol=1e-6)      # three iterations ran on what the block wrote
    assert s.domain.iter == 13
    rows = open(tmp_path / "output" / "r_probe.csv").read().split("\n")
    assert rows[0] == '"it","umax","n"'
    it, umax, n = rows[1].split(",")
    assert float(it) == 10 and float(umax) >= 0 and int(float(n)) == 32 * 16
    s.close()
