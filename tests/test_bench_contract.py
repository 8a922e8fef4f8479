"""bench.py contract pieces that can be checked without a GPU: the reference arm answers with one JSON line and exit code 0
when it cannot run (no GPU here), the weak-scaling shapes follow the reference's decomposition rule, the named
configurations are the ones of BASELINE.json."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_reports_unavailable_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("a GPU is present: the arm runs for real (covered by the driver)")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "1"],
                       cwd=ROOT, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    line = [l for l in r.stdout.splitlines() if l.strip().startswith("{")][-1]
    d = json.loads(line)
    assert d["impl"] == "reference" and isinstance(d.get("unavailable"), str) and d["unavailable"]


def test_named_configurations_and_weak_shapes():
    sys.path.insert(0, ROOT)
    import importlib
    bench = importlib.import_module("bench")
    assert set(bench.CONFIGS) == {"cumulant", "d3q19", "pf_velocity", "weak"}
    assert bench.CONFIGS["cumulant"]["shape"] == (1024, 512, 512) and bench.CONFIGS["cumulant"]["globals"]
    base = json.load(open(os.path.join(ROOT, "BASELINE.json")))
    text = json.dumps(base)
    assert "d3q27_cumulant" in text and "1024" in text
    for n, (dy, dz) in {1: (1, 1), 2: (2, 1), 4: (2, 2), 8: (4, 2)}.items():
        assert bench.weak_shape(256, n) == (256, 256 * dy, 256 * dz)
