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
