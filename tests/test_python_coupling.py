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
