"""Checkpoint images written by the device pack kernel (csrc/kernels/pri_pack.cu, streamed through a fixed staging buffer)
are byte-identical to the host packer's and load back exactly, also when the staging buffer is far smaller than a piece."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("model,shape,nranks", [("d2q9", (64, 40, 1), 2), ("d3q19", (32, 20, 16), 1), ("d3q27_cumulant", (128, 24, 20), 4)])
def test_device_packed_pri_equals_host_packed(tmp_path, monkeypatch, model, shape, nranks):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from helpers import setup_case
    from tclb_b200.runtime import Domain, Model
    m = Model(model, "float")
    g = Domain(m, shape, "cuda", nranks=nranks)
    setup_case(g, m, shape)
    g.iterate(7)
    monkeypatch.setenv("TCLB_PRI_CHUNK_MB", "1")
    for r in range(nranks):
        g.save_pri(r, str(tmp_path / f"gpu_{r}.pri"))
    c = Domain(m, shape, "cpu", nranks=nranks)
    setup_case(c, m, shape)
    for f in m.fields:
        c.set_field(f, g.get_field(f))
    for r in range(nranks):
        c.save_pri(r, str(tmp_path / f"cpu_{r}.pri"))
        a, b = open(tmp_path / f"gpu_{r}.pri", "rb").read(), open(tmp_path / f"cpu_{r}.pri", "rb").read()
        assert len(a) == len(b) > 0 and a == b
    # load on the device (again through 1 MB chunks) into a fresh domain and write once more
    h = Domain(m, shape, "cuda", nranks=nranks)
    setup_case(h, m, shape)
    for r in range(nranks):
        h.load_pri(r, str(tmp_path / f"cpu_{r}.pri"))
    for f in m.fields:
        assert np.array_equal(h.get_field(f), g.get_field(f)), f
    with open(tmp_path / "short.pri", "wb") as fh:
        fh.write(open(tmp_path / "cpu_0.pri", "rb").read()[:-8])
    with pytest.raises(Exception, match="size mismatch"):
        h.load_pri(0, str(tmp_path / "short.pri"))
    for d in (g, c, h):
        d.close()
