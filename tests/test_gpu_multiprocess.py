"""Case files on several GPUs, one process per GPU (`python -m torch.distributed.run -m tclb_b200.run --gpu ...`): the
reference's karman_p2 / karman_rs_p2 expectations (tests/external/d2q9/karman_p2.test, karman_rs_p2.test) with the halo
and Globals travelling through CUDA-IPC peer memory inside the stage kernels."""
import csv
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _launch(tmp_path, case, port, n, extra=()):
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr", "127.0.0.1",
                        "--master-port", str(port), "-m", "tclb_b200.run", "--gpu", "--precision", "double", *extra, "d2q9", case],
                       cwd=tmp_path, capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.parametrize("halo", ["direct", "nccl"])
def test_karman_case_file_on_two_gpus(tmp_path, halo):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("single GPU box")
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from golden_karman import KARMAN_IT1000, KARMAN_XML
    xml = KARMAN_XML.replace("<Solve", '<VTK Iterations="1000"/>\n<SaveCheckpoint Iterations="500" keep="all"/>\n<Solve', 1)
    (tmp_path / "karman.xml").write_text(xml)
    _launch(tmp_path, "karman.xml", 29631, 2, ("--halo", halo))
    out = tmp_path / "output"
    rows = list(csv.reader(open(out / "karman_Log_P00_00000000.csv")))
    last = dict(zip(rows[0], rows[-1]))
    for k, v in KARMAN_IT1000.items():
        assert abs(float(last[k]) - v) <= 1e-8 * abs(v), (k, last[k], v)
    names = sorted(os.listdir(out))
    assert "karman_VTK_P00_00001000.pvti" in names and "karman_VTK_P01_00001000.vti" in names
    assert "karman_checkpoint_00000500_1.pri" in names and "karman_restart_00000500.xml" in names
    _launch(tmp_path, os.path.join("output", "karman_restart_00000500.xml"), 29632, 2, ("--halo", halo))
    logs = [f for f in os.listdir(out) if f.startswith("karman_restart_00000500_Log")]
    rows2 = list(csv.reader(open(out / logs[0])))
    at500 = [dict(zip(rows2[0], r)) for r in rows2[1:] if int(float(r[0])) == 500][0]
    for k, v in KARMAN_IT1000.items():
        assert abs(float(at500[k]) - v) <= 1e-8 * abs(v), (k, at500[k], v)
