"""Multi-GPU halo exchange over NCCL, one process per GPU (needs >= 2 GPUs: run under
`gpurun --gpus N`; skipped on a single-GPU box, where the same pack/unpack/overlap logic is covered by
the in-process ranks of test_gpu_parity.py).  Bar (north_star): bit-exact against a single-GPU run of
the same global grid, and against the oracle's N-rank run."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import PHYS, ROOT

pytestmark = pytest.mark.gpu


def n_gpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("nranks", [2, 4, 8])
@pytest.mark.parametrize("prec,noise,L", [("Float32", 0.1, 320), ("Float64", 0.0, 96)])
def test_nccl_halo_bit_exact_vs_single_gpu(sim, orc, tmp_path, nranks, prec, noise, L):
    if n_gpus() < nranks:
        pytest.skip(f"needs {nranks} GPUs")
    steps = 12
    port = 29600 + (os.getpid() + nranks) % 1500
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nranks}",
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "tests", "_multi_worker.py"), str(L), prec, str(noise), str(steps), str(tmp_path)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    T = np.float32 if prec == "Float32" else np.float64
    U, V = np.empty((L, L, L), T), np.empty((L, L, L), T)
    for rk in range(nranks):
        meta = np.load(tmp_path / f"meta{rk}.npy")
        ox, oy, oz, sx, sy, sz = (int(x) for x in meta[:6])
        U[oz:oz + sz, oy:oy + sy, ox:ox + sx] = np.load(tmp_path / f"u{rk}.npy")
        V[oz:oz + sz, oy:oy + sy, ox:ox + sx] = np.load(tmp_path / f"v{rk}.npy")
    # single GPU, same global grid
    s = sim.Settings(L=L, precision=prec, backend="B200", noise=noise, **PHYS)
    mcd = sim.init_domain(s, 0, 1)
    one = sim.init_fields(s, mcd, prec, seed=9)
    sim.iterate(one, s, mcd, nsteps=steps)
    u1, v1 = sim.get_fields(one)
    one.close()
    assert np.array_equal(U, u1) and np.array_equal(V, v1)
    if L <= 128:
        o = orc.Oracle(L, prec, dims=orc.dims_create(nranks), noise=noise, seed=9, **PHYS)
        o.step(steps)
        uo, vo = o.gather()
        assert np.array_equal(U, uo) and np.array_equal(V, vo)


@pytest.mark.parametrize("prec,noise", [("Float32", 0.1), ("Float64", 0.0)])
@pytest.mark.parametrize("scheme", ["pipeline", "split"])
def test_inprocess_ranks_on_separate_gpus(sim, orc, monkeypatch, prec, noise, scheme):
    """gsb_comm_init_local with the handles on DIFFERENT devices of one process: the pack kernel stores
    straight into the neighbour GPU's receive buffer over NVLink (peer access), events order the two
    devices.  Must equal the oracle's N-rank run bit for bit."""
    ndev = n_gpus()
    if ndev < 2:
        pytest.skip("needs >= 2 GPUs")
    monkeypatch.setenv("GSB_HALO", scheme)
    nranks = 8 if ndev >= 8 else (4 if ndev >= 4 else 2)
    L, steps = 128, 9
    s = sim.Settings(L=L, precision=prec, backend="B200", noise=noise, **PHYS)
    ranks = []
    for r in range(nranks):
        mcd = sim.init_domain(s, r, nranks)
        ranks.append((mcd, sim.init_fields(s, mcd, prec, seed=9, device=r % ndev)))
    sim.connect_local([f for _, f in ranks])
    sim.iterate_group([f for _, f in ranks], s, nsteps=steps)
    o = orc.Oracle(L, prec, dims=orc.dims_create(nranks), noise=noise, seed=9, **PHYS)
    o.step(steps)
    for r, (mcd, f) in enumerate(ranks):
        u, v = sim.get_fields(f)
        uo, vo = o.get_fields(r)
        assert np.array_equal(u, uo) and np.array_equal(v, vo), r
        f.close()
