"""<HDF5> element: .h5 written by the built-in minimal writer + .xmf index (reference src/hdf5Lattice.cpp)."""
import os
import re
import subprocess

import numpy as np

from golden_karman import KARMAN_XML


def test_hdf5_output_matches_vtk_fields(tclb_bin, tmp_path):
    from tclb_b200 import build
    from tclb_b200.utils import h5min
    build.build_model_lib("d2q9", "double")
    xml = KARMAN_XML.replace('<Solve Iterations="1000"', '<HDF5 Iterations="50" what="U,Rho,BOUNDARY" precision="double"/><TXT Iterations="50" what="Rho"/><Solve Iterations="50"')
    assert "<HDF5" in xml
    (tmp_path / "k.xml").write_text(xml)
    r = subprocess.run([tclb_bin, "--cpu", "--precision", "double", "--ranks", "2", "d2q9", "k.xml"], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    h5 = tmp_path / "output" / "k_HDF5_00000050.h5"
    xmf = tmp_path / "output" / "k_HDF5_00000050.xmf"
    assert h5.exists() and xmf.exists(), os.listdir(tmp_path / "output")
    d = h5min.read(str(h5))
    assert set(d) == {"U", "Rho", "BOUNDARY"}
    nz, ny, nx = d["Rho"].shape
    assert (nz, ny, nx) == (1, 100, 1024) and d["U"].shape == (1, 100, 1024, 3) and d["BOUNDARY"].dtype == np.uint8
    assert np.isfinite(d["U"]).all() and 0.9 < d["Rho"].mean() < 1.1 and d["BOUNDARY"].max() > 0
    txt = np.loadtxt(tmp_path / "output" / "k_TXT_P00_00000050_Rho.txt")
    assert txt.size > 0
    x = open(xmf).read()
    assert 'Type="3DCoRectMesh"' in x and "k_HDF5_00000050.h5:/U" in x and re.search(r'Dimensions="1 100 1024 3"', x)


def test_hdf5_pieces_from_two_processes(tmp_path):
    """One process per subdomain: every process writes its piece, rank 0 the XDMF spatial collection; the pieces put together
    equal the single-file output of the one-process run (reference: collective write, src/hdf5Lattice.cpp:154-195)."""
    import sys
    from tclb_b200 import build
    from tclb_b200.utils import h5min
    ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    build.build_model_lib("d2q9", "double")
    xml = KARMAN_XML.replace('<Solve Iterations="1000"', '<HDF5 Iterations="50" what="U,Rho,BOUNDARY" precision="double"/><Solve Iterations="50"')
    one, two = tmp_path / "one", tmp_path / "two"
    one.mkdir(); two.mkdir()
    (one / "k.xml").write_text(xml); (two / "k.xml").write_text(xml)
    env = dict(os.environ, OMP_NUM_THREADS="2", PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    r = subprocess.run([sys.executable, "-m", "tclb_b200.run", "--cpu", "--precision", "double", "--ranks", "2", "d2q9", "k.xml"], cwd=one, capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1", "--master-port", "29641",
                        "-m", "tclb_b200.run", "--cpu", "--precision", "double", "d2q9", "k.xml"], cwd=two, capture_output=True, text=True, env=env, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    whole = h5min.read(str(one / "output" / "k_HDF5_00000050.h5"))
    p0 = h5min.read(str(two / "output" / "k_HDF5_P00_00000050.h5"))
    p1 = h5min.read(str(two / "output" / "k_HDF5_P01_00000050.h5"))
    for name in ("U", "Rho", "BOUNDARY"):
        assert np.array_equal(np.concatenate([p0[name], p1[name]], axis=1), whole[name]), name
    x = open(two / "output" / "k_HDF5_00000050.xmf").read()
    assert 'CollectionType="Spatial"' in x and "k_HDF5_P00_00000050.h5:/U" in x and "k_HDF5_P01_00000050.h5:/Rho" in x
    assert x.count("<Grid Name=\"Lattice_P") == 2 and x.count('Dimensions="1 50 1024 3"') == 2


def test_hdf5_compressed_output_equals_contiguous(tclb_bin, tmp_path):
    """<HDF5 compress=...>: chunked datasets (slabs along z, version-1 chunk B-tree) with the deflate filter carry the same
    arrays as the contiguous file and are smaller; a 3-d case with more slabs than one index node holds uses two levels"""
    from tclb_b200 import build
    from tclb_b200.utils import h5min
    build.build_model_lib("d2q9", "double")
    build.build_model_lib("d3q19", "double")
    xml = KARMAN_XML.replace('<Solve Iterations="1000"', '<HDF5 Iterations="50" name="plain" what="U,Rho,BOUNDARY" precision="double"/>'
                             '<HDF5 Iterations="50" name="packed" what="U,Rho,BOUNDARY" precision="double" compress="true"/><Solve Iterations="50"')
    (tmp_path / "k.xml").write_text(xml)
    r = subprocess.run([tclb_bin, "--cpu", "--precision", "double", "d2q9", "k.xml"], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    a = h5min.read(str(tmp_path / "output" / "k_plain_00000050.h5"))
    p = tmp_path / "output" / "k_packed_00000050.h5"
    b_ = h5min.read(str(p))
    assert set(a) == set(b_)
    for k in a:
        assert a[k].dtype == b_[k].dtype and np.array_equal(a[k], b_[k]), k
    assert os.path.getsize(p) < 0.8 * os.path.getsize(tmp_path / "output" / "k_plain_00000050.h5")
    # 3-d, 70 slabs: one internal node over two leaves
    xml3 = """<?xml version="1.0"?>
<CLBConfig version="2.0" output="output/">
  <Geometry nx="16" ny="12" nz="70"><MRT><Box/></MRT><Wall mask="ALL"><Box ny="1"/><Box dy="-1"/></Wall></Geometry>
  <Model><Param name="nu" value="0.05"/><Param name="ForceX" value="1e-5"/></Model>
  <HDF5 Iterations="10" name="plain" what="U,Rho"/>
  <HDF5 Iterations="10" name="packed" what="U,Rho" compress="4"/>
  <Solve Iterations="10"/>
</CLBConfig>
"""
    (tmp_path / "c.xml").write_text(xml3)
    r = subprocess.run([tclb_bin, "--cpu", "--precision", "double", "d3q19", "c.xml"], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    a = h5min.read(str(tmp_path / "output" / "c_plain_00000010.h5"))
    b_ = h5min.read(str(tmp_path / "output" / "c_packed_00000010.h5"))
    assert a["U"].shape == (70, 12, 16, 3) and a["U"].dtype == np.float32
    for k in a:
        assert np.array_equal(a[k], b_[k]), k
