"""tclb_b200.ops: the runtime's stand-alone kernels on torch tensors, against plain PyTorch (fp32/fp64 references)."""
import numpy as np
import pytest
import torch

from tclb_b200 import ops


def _rows(n, dtype, device, seed=0):
    g = torch.Generator().manual_seed(seed)
    r = torch.zeros((n, 16), dtype=dtype)
    r[:, 0] = 0.4 + 0.8 * torch.rand(n, generator=g, dtype=dtype)
    r[:, 1:4] = 32.0 * torch.rand((n, 3), generator=g, dtype=dtype)
    return r.to(device)


def _grid_reference(rows, origin, delta, cells, depth):
    cx, cy, cz = cells
    want = [[set() for _ in range(cx * cy * cz)]]
    r = rows.cpu().double().numpy()
    sets = want[0]
    for i in range(r.shape[0]):
        c = np.floor((r[i, 1:4] - np.array(origin)) / delta).astype(int)
        if all(0 <= c[k] < (cx, cy, cz)[k] for k in range(3)):
            sets[(c[2] * cy + c[1]) * cx + c[0]].add(i)
    return sets


def _check_grid(device):
    rows = _rows(200, torch.float64, device)
    cells, delta, origin, depth = (8, 8, 8), 4.0, (0.0, 0.0, 0.0), 16
    grid, counts, overflow = ops.particle_grid(rows, origin, delta, cells, depth)
    assert int(overflow.cpu()) == 0
    sets = _grid_reference(rows, origin, delta, cells, depth)
    g = grid.cpu().numpy().reshape(-1, depth)
    c = counts.cpu().numpy().reshape(-1)
    for k, s in enumerate(sets):
        assert c[k] == len(s) and set(int(v) for v in g[k] if v >= 0) == s, k


def test_particle_grid_cpu(built):
    _check_grid("cpu")


def test_accumulate_nonzero_cpu(built):
    a = torch.randn(1000, dtype=torch.float64)
    b = torch.randn(1000, dtype=torch.float64) * (torch.rand(1000) > 0.5)
    want = a + b
    assert torch.equal(ops.accumulate_nonzero(a, b), want)


@pytest.mark.gpu
def test_particle_grid_gpu():
    _check_grid("cuda:0")


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_scan_and_accumulate_gpu(dtype):
    t = torch.randn(3_000_001, dtype=dtype, device="cuda:0")
    t[12345] = float("nan")
    for static in (False, True):
        got = ops.scan(t, static=static).cpu().tolist()
        ref = t.double()
        assert got[0] == 1.0 and got[1] == float(ref[~ref.isnan()].min()) and got[2] == float(ref[~ref.isnan()].max())
    a = torch.randn(1 << 20, dtype=dtype, device="cuda:0")
    b = torch.randn(1 << 20, dtype=dtype, device="cuda:0") * (torch.rand(1 << 20, device="cuda:0") > 0.5)
    want = (a.double() + b.double()).to(dtype)
    assert torch.equal(ops.accumulate_nonzero(a, b), want)
