"""Zero-copy coupling with torch (north star: lattice fields exposed as torch tensors via DLPack): a tensor obtained from
`LatticeField.__dlpack__` / `interop.field_view` is the lattice memory itself - a write through the tensor is what
`get_field` returns and what the next iteration streams."""
import numpy as np
import pytest

from helpers import setup_case


def _domain(backend):
    from tclb_b200.runtime import Domain, Model
    m = Model("d2q9", "double")
    d = Domain(m, (32, 16, 1), backend, nranks=1)
    setup_case(d, m, (32, 16, 1))
    d.iterate(5)
    return m, d


def _check(backend):
    import torch
    from tclb_b200.utils.interop import LatticeField, field_view
    m, d = _domain(backend)
    name = m.fields[3]
    ref = d.get_field(name)
    t = torch.from_dlpack(LatticeField(d, name))
    assert tuple(t.shape) == ref.shape and t.dtype == torch.float64
    assert np.array_equal(t.cpu().numpy(), ref)
    v = field_view(d, name)
    assert v.data_ptr() == t.data_ptr() and v.stride() == t.stride()       # both are the same memory, not copies
    t[0, 5, 7] = 0.125                                                        # write through the DLPack tensor ...
    if backend == "cuda":
        torch.cuda.synchronize()
    assert d.get_field(name)[0, 5, 7] == 0.125                               # ... is seen by the lattice
    before = d.get_quantity("Rho")
    d.iterate(1)
    after = d.get_quantity("Rho")
    assert np.abs(after - before).max() > 1e-3                               # and by the next iteration
    d.close()


def test_dlpack_field_is_lattice_memory_cpu():
    _check("cpu")


@pytest.mark.gpu
def test_dlpack_field_is_lattice_memory_gpu():
    _check("cuda")
