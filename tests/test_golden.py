"""Committed golden vectors (tests/golden/*.npz, written by tests/golden/make_golden.py): the oracle must
still reproduce them byte for byte (CPU), and so must the CUDA path through the C ABI (GPU)."""
import glob
import os

import numpy as np
import pytest

from conftest import PHYS, ROOT

FILES = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "*.npz")))


def load(path):
    d = np.load(path)
    return d, int(d["L"]), str(d["precision"]), int(d["steps"]), float(d["noise"]), int(d["seed"]), tuple(int(x) for x in d["dims"])


def test_golden_files_present():
    assert len(FILES) >= 5


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f)[:-4] for f in FILES])
def test_oracle_reproduces_golden(orc, path):
    d, L, prec, steps, noise, seed, dims = load(path)
    o = orc.Oracle(L, prec, dims=dims, noise=noise, seed=seed, **PHYS)
    o.step(steps)
    u, v = o.gather()
    assert u.tobytes() == d["u"].tobytes() and v.tobytes() == d["v"].tobytes()


@pytest.mark.gpu
@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f)[:-4] for f in FILES])
def test_cuda_path_reproduces_golden(sim, path):
    d, L, prec, steps, noise, seed, dims = load(path)
    nranks = dims[0] * dims[1] * dims[2]
    s = sim.Settings(L=L, precision=prec, backend="B200", noise=noise, **PHYS)
    ranks = []
    for r in range(nranks):
        mcd = sim.init_domain(s, r, nranks)
        assert mcd.dims == dims
        ranks.append((mcd, sim.init_fields(s, mcd, prec, seed=seed)))
    if nranks > 1:
        sim.connect_local([f for _, f in ranks])
    sim.iterate_group([f for _, f in ranks], s, nsteps=steps)
    U, V = np.empty_like(d["u"]), np.empty_like(d["v"])
    for mcd, f in ranks:
        u, v = sim.get_fields(f)
        ox, oy, oz = mcd.proc_offsets
        sx, sy, sz = mcd.proc_sizes
        U[oz:oz + sz, oy:oy + sy, ox:ox + sx] = u
        V[oz:oz + sz, oy:oy + sy, ox:ox + sx] = v
        f.close()
    assert U.tobytes() == d["u"].tobytes() and V.tobytes() == d["v"].tobytes()
