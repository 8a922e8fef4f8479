"""End-to-end regression against the reference's own golden log (SURVEY.md section 4:
tests/external/d2q9/karman.test, karman_p2.test, karman_rs.test)."""
import csv
import os
import subprocess

import pytest

from golden_karman import KARMAN_IT1000, KARMAN_XML


def _run(tclb_bin, tmp_path, args, xml=KARMAN_XML, name="karman.xml", model="d2q9"):
    (tmp_path / name).write_text(xml)
    cmd = [tclb_bin, "--cpu", "--precision", "double"] + args + [model, name]
    r = subprocess.run(cmd, cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    return r


def _log(path):
    rows = list(csv.reader(open(path)))
    return rows[0], rows[1:]


def _check_it1000(hdr, rows, tol=1e-9):
    last = dict(zip(hdr, rows[-1]))
    assert int(last["Iteration"]) == 1000
    for k, v in KARMAN_IT1000.items():
        assert abs(float(last[k]) - v) <= tol * abs(v), (k, last[k], v)


def test_karman_matches_reference_golden(tclb_bin, tmp_path):
    from tclb_b200 import build
    build.build_model_lib("d2q9", "double")
    _run(tclb_bin, tmp_path, [])
    hdr, rows = _log(tmp_path / "output" / "karman_Log_P00_00000000.csv")
    assert len(rows) == 10
    _check_it1000(hdr, rows)
    assert (tmp_path / "output" / "karman_config_P00_00000000.xml").exists()


def test_karman_full_log_vs_mounted_reference(tclb_bin, tmp_path, reference_dir):
    from tclb_b200 import build
    build.build_model_lib("d2q9", "double")
    _run(tclb_bin, tmp_path, [])
    hdr, rows = _log(tmp_path / "output" / "karman_Log_P00_00000000.csv")
    ghdr, grows = _log(os.path.join(reference_dir, "tests/external/d2q9/output/karman_Log_P00_00000000.csv"))
    assert hdr == ghdr
    for r, g in zip(rows, grows):
        for n, a, b in zip(hdr, r, g):
            if n == "Walltime":
                continue
            a, b = float(a), float(b)
            assert abs(a - b) <= 1e-9 * max(abs(b), 1e-300) or abs(a - b) < 1e-300, (n, a, b)


@pytest.mark.parametrize("ranks", [2, 4])
def test_karman_decomposition_invariance(tclb_bin, tmp_path, ranks):
    """same goldens with the lattice split over several subdomains (reference karman_p2.test)"""
    from tclb_b200 import build
    build.build_model_lib("d2q9", "double")
    _run(tclb_bin, tmp_path, ["--ranks", str(ranks)])
    hdr, rows = _log(tmp_path / "output" / "karman_Log_P00_00000000.csv")
    _check_it1000(hdr, rows)


def test_karman_restart(tclb_bin, tmp_path):
    """run 500 + restart 500 == 1000 (reference karman_rs.test); checkpoint + restart files exist"""
    from tclb_b200 import build
    build.build_model_lib("d2q9", "double")
    xml = KARMAN_XML.replace('<Solve Iterations="1000"><Log Iterations="100"/></Solve>',
                             '<Solve Iterations="500"><Log Iterations="100"/></Solve><SaveCheckpoint/>')
    _run(tclb_bin, tmp_path, [], xml=xml)
    pri = tmp_path / "output" / "karman_checkpoint_00000500_0.pri"
    rst = tmp_path / "output" / "karman_restart_00000500.xml"
    assert pri.exists() and rst.exists()
    text = rst.read_text()
    assert "LoadBinary" in text
    # resume: 500 more iterations from the restart file
    r = subprocess.run([tclb_bin, "--cpu", "--precision", "double", "d2q9", str(rst),
                        "/CLBConfig/SaveCheckpoint", "delete"], cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    logs = sorted((tmp_path / "output").glob("karman_restart_00000500_Log_*.csv"))
    assert logs, os.listdir(tmp_path / "output")
    hdr, rows = _log(logs[0])
    last = dict(zip(hdr, rows[-1]))
    # the resumed run counts iterations from 0, its 500th iteration is the 1000th overall
    assert int(last["Iteration"]) == 500
    for k, v in KARMAN_IT1000.items():
        assert abs(float(last[k]) - v) <= 1e-9 * abs(v), (k, last[k], v)


def test_bad_input_fails(tclb_bin, tmp_path):
    """reference fails.test: malformed cases must fail"""
    (tmp_path / "bad.xml").write_text("<CLBConfig><Geometry nx='32' ny='8'><NoSuchType><Box/></NoSuchType></Geometry><Model/></CLBConfig>")
    r = subprocess.run([tclb_bin, "--cpu", "--precision", "double", "d2q9", "bad.xml"], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode != 0
    (tmp_path / "bad2.xml").write_text("<CLBConfig><Geometry nx='32' ny='8'></CLBConfig>")
    r = subprocess.run([tclb_bin, "--cpu", "--precision", "double", "d2q9", "bad2.xml"], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode != 0


def test_image_element_writes_the_colour_kernel_picture(tclb_bin, tmp_path):
    """<Image>: the headless counterpart of the reference's interactive view (velocity magnitude through the colour map,
    src/LatticeContainer.inc.cpp.Rt:355-423): one binary PPM per call, equal to Domain.color() of the same state"""
    import subprocess
    import numpy as np
    from golden_karman import KARMAN_XML
    xml = KARMAN_XML.replace('<Solve Iterations="1000"', '<Image Iterations="40" scale="20"/><Solve Iterations="40"')
    (tmp_path / "k.xml").write_text(xml)
    r = subprocess.run([tclb_bin, "--cpu", "--precision", "double", "--ranks", "2", "d2q9", "k.xml"], cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    raw = open(tmp_path / "output" / "k_Image_00000040.ppm", "rb").read()
    head, w_h, mx, rest = raw.split(b"\n", 3)
    assert head == b"P6" and w_h == b"1024 100" and mx == b"255" and len(rest) == 1024 * 100 * 3
    img = np.frombuffer(rest, dtype=np.uint8).reshape(100, 1024, 3)
    assert len(np.unique(img.reshape(-1, 3), axis=0)) > 20          # a picture, not a constant
    assert not np.array_equal(img[50, 100:200], img[50, 600:700])   # wake behind the obstacle differs from the inflow


def test_case_with_units_element_runs_from_the_command_line(tclb_bin, tmp_path):
    """a case in physical units (the reference's example/flow/2d/D2Q9_Poiseuille.xml pattern): <Units> is read before the
    lattice exists; the gauge is printed and lengths / viscosities are converted"""
    import subprocess
    xml = """<?xml version="1.0"?>
<CLBConfig version="2.0" output="output/" permissive="true">
  <Units>
    <Param name="size" value="0.0005m" gauge="1"/>
    <Param name="nu" value="1e-5m2/s" gauge="0.1666666666"/>
  </Units>
  <Geometry nx="0.02m" ny="0.0105m">
    <MRT><Box/></MRT>
    <Wall mask="ALL"><Channel/></Wall>
  </Geometry>
  <Model>
    <Param name="Viscosity" value="1e-5m2/s"/>
    <Param name="GravitationX" value="1e-4m/s2"/>
  </Model>
  <Log Iterations="5"/>
  <Solve Iterations="10"/>
</CLBConfig>
"""
    (tmp_path / "u.xml").write_text(xml)
    r = subprocess.run([tclb_bin, "--cpu", "--precision", "double", "d2q9", "u.xml"], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    out = r.stdout + r.stderr
    assert r.returncode == 0, out[-2000:]
    assert "GAUGE" in out and "Global lattice size: 40x21x1" in out
    assert "0.1666" in out and "1 s = 240.000000 units" in out          # 1e-5 m2/s in lattice units; the time scale that follows
