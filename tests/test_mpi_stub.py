"""The single-node MPI stand-in the reference arm is linked against (baseline/mpi_stub): its self-test as 1, 2 and 4
processes - collectives, and the reference's halo pattern (Irecv/Isend/Waitany with several tags per pair and messages
larger than one ring)."""
import os
import random
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STUB = os.path.join(ROOT, "baseline", "mpi_stub")


@pytest.fixture(scope="module")
def selftest(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("mpi") / "selftest")
    subprocess.run(["g++", "-O2", "-std=c++17", "-I", STUB, os.path.join(STUB, "test", "selftest.cpp"), os.path.join(STUB, "mpi_stub.cpp"),
                    "-o", exe, "-lpthread"], check=True)
    return exe


@pytest.mark.parametrize("n", [1, 2, 4])
def test_mpi_stub_selftest(selftest, n, tmp_path):
    job = str(random.randrange(1 << 30))
    procs = [subprocess.Popen([selftest], env=dict(os.environ, TCLB_MPI_RANK=str(r), TCLB_MPI_SIZE=str(n), TCLB_MPI_ID=job, TCLB_MPI_DIR=str(tmp_path)),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(n)]
    outs = [p.communicate(timeout=300)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert f"selftest ok on {n} ranks" in outs[0]
