"""GPU tests (run on the B200 box with `-m gpu`): the sm_100a kernels against the CPU restatement of
the same generated node code, fused halo vs copy baseline on several subdomains, golden karman log."""
import csv
import os
import subprocess

import numpy as np
import pytest

from helpers import run_case

pytestmark = pytest.mark.gpu

CASES = [("d2q9", (128, 48, 1)), ("d3q19", (64, 24, 20)), ("d3q27_cumulant", (64, 24, 20))]


@pytest.fixture(scope="module", autouse=True)
def libs(built):
    from tclb_b200 import build
    for m, _ in CASES:
        for p in ("float", "double"):
            build.build_model_lib(m, p)
    from tclb_b200.runtime import core
    core()          # fails loudly if the native library is missing


@pytest.mark.parametrize("model,shape", CASES)
def test_gpu_double_matches_cpu_oracle(model, shape):
    a = run_case(model, "double", "cpu", shape, 40)
    b = run_case(model, "double", "cuda", shape, 40)
    for q in a:
        if q.startswith("_"):
            continue
        assert np.abs(a[q] - b[q]).max() < 1e-12, (model, q, np.abs(a[q] - b[q]).max())   # FMA contraction only
    for k, v in a["_globals"].items():
        assert abs(v - b["_globals"][k]) <= 1e-10 * max(1.0, abs(v)), (k, v, b["_globals"][k])


@pytest.mark.parametrize("model,shape", CASES)
def test_gpu_float_matches_fp32_reference(model, shape):
    a = run_case(model, "double", "cpu", shape, 40)
    b = run_case(model, "float", "cuda", shape, 40)
    assert np.abs(a["U"] - b["U"]).max() < 5e-6
    n = float(np.prod(shape))
    for k, v in a["_globals"].items():
        # fp32 node values carry ~1e-7 relative rounding; a sum over n nodes of O(1) populations may drift by n*1e-7
        assert abs(v - b["_globals"][k]) <= 2e-4 * max(1.0, abs(v)) + 2e-7 * n, (k, v, b["_globals"][k])


@pytest.mark.parametrize("model,shape", CASES)
def test_gpu_virtual_ranks_fused_halo(model, shape):
    """4 subdomains on one GPU: fused in-kernel ghost stores == memcpy exchange == single subdomain (bitwise)"""
    a = run_case(model, "float", "cuda", shape, 25, nranks=1)
    b = run_case(model, "float", "cuda", shape, 25, nranks=4, direct=True)
    c = run_case(model, "float", "cuda", shape, 25, nranks=4, direct=False)
    for q in a:
        if q.startswith("_"):
            continue
        assert np.array_equal(a[q], b[q]), (model, q)
        assert np.array_equal(b[q], c[q]), (model, q)


def test_gpu_karman_golden(tclb_bin, tmp_path):
    from golden_karman import KARMAN_IT1000, KARMAN_XML
    (tmp_path / "karman.xml").write_text(KARMAN_XML)
    r = subprocess.run([tclb_bin, "--gpu", "--precision", "double", "d2q9", "karman.xml"], cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    rows = list(csv.reader(open(tmp_path / "output" / "karman_Log_P00_00000000.csv")))
    last = dict(zip(rows[0], rows[-1]))
    for k, v in KARMAN_IT1000.items():
        assert abs(float(last[k]) - v) <= 1e-8 * abs(v), (k, last[k], v)


def test_gpu_smoke_entry():
    import __graft_entry__ as g
    g.smoke()


def test_multi_gpu_direct_halo_if_available():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("single GPU box")
    n = min(4, torch.cuda.device_count())
    a = run_case("d3q27_cumulant", "float", "cuda", (64, 32, 32), 30, nranks=1)
    b = run_case("d3q27_cumulant", "float", "cuda", (64, 32, 32), 30, nranks=n, devices=list(range(n)))
    assert np.array_equal(a["U"], b["U"])


def _scan(lib, t):
    import ctypes as C
    import torch
    out = torch.zeros(3, dtype=torch.float64, device=t.device)
    lib.tclb_scan_array.argtypes = [C.c_void_p, C.c_longlong, C.c_int, C.c_void_p, C.c_void_p]
    lib.tclb_scan_last_error.restype = C.c_char_p
    rc = lib.tclb_scan_array(t.data_ptr(), t.numel(), t.element_size(), None, out.data_ptr())
    assert rc == 0, lib.tclb_scan_last_error()
    torch.cuda.synchronize()
    return out.cpu().numpy()


@pytest.mark.parametrize("n,dtype", [(1, "float32"), (4097, "float32"), (1 << 20, "float64"), (3_000_001, "float32")])
def test_device_scan_matches_numpy(n, dtype):
    import torch
    from tclb_b200.runtime import core
    lib = core()
    rng = np.random.default_rng(n)
    a = rng.standard_normal(n).astype(dtype)
    k = min(n, 5)
    idx = rng.choice(n, size=k, replace=False) if n > 5 else np.array([], dtype=int)
    a[idx] = np.nan
    t = torch.from_numpy(a).cuda()
    for _ in range(2):
        r = _scan(lib, t)
        fin = a[np.isfinite(a)]
        assert r[0] == len(idx)
        if len(fin):
            assert r[1] == float(fin.min()) and r[2] == float(fin.max())


def test_failcheck_stops_on_nan(tclb_bin, tmp_path):
    xml = """<CLBConfig output="output/"><Geometry nx="64" ny="32"><MRT><Box/></MRT><Wall mask="ALL"><Channel/></Wall></Geometry>
    <Model><Param name="Viscosity" value="-10"/><Param name="GravitationX" value="0.3"/></Model>
    <Failcheck Iterations="20"/><Solve Iterations="4000"/></CLBConfig>"""
    (tmp_path / "nan.xml").write_text(xml)
    r = subprocess.run([tclb_bin, "--gpu", "d2q9", "nan.xml"], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "NaN" in r.stdout


def test_gpu_globals_reproducible_and_objective_weights():
    """Per-block partials + fixed-order fold: two identical runs give bitwise identical Globals; the Objective
    terms follow the <Global>InObj zone weights and vanish (skipped in-kernel) when all weights are zero."""
    from helpers import setup_case
    from tclb_b200.runtime import Domain, Model
    m = Model("d3q27_cumulant", "float")
    shape = (96, 40, 24)
    res = []
    for _ in range(2):
        d = Domain(m, shape, "cuda", nranks=1)
        setup_case(d, m, shape)
        d.iterate(12, globals_every=True)
        res.append(d.globals())
    assert res[0] == res[1]
    assert res[0]["Objective"] == 0.0
    fluid = int((d.get_flags() & m.nodetypes["MRT"][1] == m.nodetypes["MRT"][0]).sum())
    d.set_setting("MassInObj", 0.5)
    d.iterate(1, globals_every=True)
    g = d.globals()
    assert abs(g["Objective"] - 0.5 * g["Mass"]) <= 1e-5 * fluid and g["Objective"] != 0.0
    md = Model("d3q27_cumulant", "double")
    cpu = Domain(md, shape, "cpu", nranks=1)
    setup_case(cpu, md, shape)
    cpu.iterate(13, globals_every=True)
    ref = cpu.globals()
    got = d.globals()
    for k in ("Mass", "XMomentum", "Flux"):
        assert abs(got[k] - ref[k]) <= 2e-4 * max(1.0, abs(ref[k])) + 2e-7 * float(np.prod(shape)), (k, got[k], ref[k])


def test_gpu_reference_pf_velocity_matches_cpu_oracle(tclb_bin, tmp_path):
    """d3q27_pf_velocity lowered from the reference's model files (library built when the reference tree is mounted):
    the fused multi-stage GPU kernels against the CPU restatement of the same generated code, through the case runner"""
    import csv
    import shutil
    import subprocess
    lib = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tclb_b200", "_lib", "libtclb_ref_d3q27_pf_velocity_double.so")
    if not os.path.exists(lib):
        pytest.skip("ref_d3q27_pf_velocity was not built (no reference tree at build time)")
    case = os.path.join(os.path.dirname(os.path.abspath(__file__)), "cases", "pf_drop_on_wall.xml")
    logs = {}
    for backend in ("cpu", "gpu"):
        wd = tmp_path / backend
        wd.mkdir()
        shutil.copy(case, wd / "case.xml")
        r = subprocess.run([tclb_bin, f"--{backend}", "--precision", "double", "ref_d3q27_pf_velocity", "case.xml"], cwd=wd,
                           capture_output=True, text=True, timeout=900)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        logs[backend] = list(csv.reader(open(wd / "output" / "case_Log_P00_00000000.csv")))
    hdr = logs["cpu"][0]
    assert hdr == logs["gpu"][0] and len(logs["cpu"]) == len(logs["gpu"]) == 7
    for a, b in zip(logs["cpu"][1:], logs["gpu"][1:]):
        for n, x, y in zip(hdr, a, b):
            if n == "Walltime":
                continue
            x, y = float(x), float(y)
            assert abs(x - y) <= 1e-9 * max(abs(x), abs(y)) + 1e-11, (n, x, y)
    liq = [float(r[hdr.index("LiqTotalPhase")]) for r in logs["gpu"][1:]]
    assert liq[0] > 100 and abs(liq[-1] - liq[0]) < 0.05 * liq[0]


def test_gpu_particles_match_cpu_oracle(tclb_bin, tmp_path):
    """particle stages on the GPU (atomic force accumulation, periodic images of a particle on the corner) against the
    CPU oracle: fluid log and particle force log of the PSM model lowered from the reference's files"""
    import csv
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if not os.path.exists(os.path.join(root, "tclb_b200", "_lib", "libtclb_ref_d3q27_PSM_NEBB_double.so")):
        pytest.skip("ref_d3q27_PSM_NEBB was not built (no reference tree at build time)")
    case = """<?xml version="1.0"?>
<CLBConfig version="2.0" permissive="true" output="output/">
  <Geometry nx="32" ny="24" nz="16"><BGK><Box/></BGK></Geometry>
  <Model>
    <Param name="nu" value="0.03"/>
    <RemoteForceInterface integrator="SIMPLEPART">
      <SimplePart dt="1" ax="0.001" afreq="0.01">
        <Particle x="1" y="12" z="0" r="4" log="y"/>
        <Particle x="20" y="6" z="9" r="3" vy="0.002" log="y"/>
        <Periodic x="32" y="24" z="16"/>
        <Log name="output/forces.csv" Iterations="1"/>
      </SimplePart>
    </RemoteForceInterface>
  </Model>
  <Log Iterations="5"/>
  <Solve Iterations="30"/>
</CLBConfig>
"""
    out = {}
    for backend in ("cpu", "gpu"):
        wd = tmp_path / backend
        wd.mkdir()
        (wd / "p.xml").write_text(case)
        r = subprocess.run([tclb_bin, f"--{backend}", "--precision", "double", "ref_d3q27_PSM_NEBB", "p.xml"], cwd=wd, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        out[backend] = [list(csv.reader(open(wd / "output" / f))) for f in ("forces.csv", "p_Log_P00_00000000.csv")]
    for a, b in zip(out["cpu"], out["gpu"]):
        assert a[0] == b[0] and len(a) == len(b) > 5
        for ra, rb in zip(a[1:], b[1:]):
            for n, x, y in zip(a[0], ra, rb):
                if n == "Walltime":
                    continue
                assert abs(float(x) - float(y)) <= 1e-9 * max(abs(float(x)), abs(float(y))) + 1e-12, (n, x, y)
    fx = [abs(float(r[8])) for r in out["gpu"][0][1:]]
    assert max(fx) > 1e-3                                   # the particles do exchange momentum with the fluid


@pytest.mark.parametrize("model,shape,nranks", [("d2q9", (64, 40, 1), 1), ("d3q19", (32, 20, 16), 2), ("d3q27_cumulant", (32, 20, 16), 4)])
def test_gpu_graph_replay_is_bitwise_identical(model, shape, nranks):
    """iterate() replays a captured two-iteration CUDA graph (iteration number, halo epoch and the time index of zonal
    series come from device memory): results equal the plain launch loop bit for bit, also after a setting changes"""
    from helpers import setup_case
    from tclb_b200.runtime import Domain, Model
    m = Model(model, "float")
    out = []
    for graphs in (True, False):
        d = Domain(m, shape, "cuda", nranks=nranks)
        d.use_graphs(graphs)
        setup_case(d, m, shape)
        zs = [z for z in m.zonesettings if not z.endswith("InObj")]
        if zs:
            d.set_zone_series(zs[0], 0, [0.01 + 0.001 * k for k in range(7)])  # time-dependent zonal setting
        d.iterate(41)
        d.set_setting(m.settings[0], d.get_setting(m.settings[0]) * 1.01)      # invalidates the captured parameters
        d.iterate(30, globals_on_last=True)
        out.append(({f: d.get_field(f) for f in m.fields[:5]}, d.globals(), d.graph_iterations, d.iter))
    assert out[0][2] >= 60 and out[1][2] == 0 and out[0][3] == out[1][3] == 71
    for f in out[0][0]:
        assert np.array_equal(out[0][0][f], out[1][0][f]), f
    assert out[0][1] == out[1][1]
