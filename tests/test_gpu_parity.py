"""GPU parity tests: the CUDA path, called through the C ABI (include/gsb200.h via ctypes), against the
CPU oracle on the same inputs.

Bars (BASELINE.json north_star): noise = 0 -> within 1e-5 relative (Float32) / 1e-12 (Float64) of the
reference CPU solver; halo exchange bit-exact against a single-subdomain run; noise on -> statistical.
GSB_ARITH_REFERENCE goes further: it must be BIT-IDENTICAL to the oracle (which restates the reference
arithmetic rounding for rounding), for both precisions.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import PHYS, PHYS_DEFAULT

pytestmark = pytest.mark.gpu

REL_TOL = {"Float32": 1e-5, "Float64": 1e-12}       # north_star tolerances
NATIVE, REFERENCE = 1, 0
KERNELS = [(0, 8, 4, 64), (0, 8, 6, 32), (0, 16, 3, 64), (0, 16, 4, 16), (1, 4, 0, 64), (1, 8, 0, 16)]  # kind, ty, stages, zchunk


def rel_err(a, b):
    """max|a-b| / max|b| : the field-normalised relative error the 1e-5 / 1e-12 bars refer to."""
    return float(np.abs(a.astype(np.float64) - b.astype(np.float64)).max() / max(np.abs(b).max(), 1e-300))


def ks_distance(a, b):
    a, b = np.sort(a.ravel().astype(np.float64)), np.sort(b.ravel().astype(np.float64))
    allv = np.concatenate([a, b])
    return float(np.abs(np.searchsorted(a, allv, side="right") / a.size - np.searchsorted(b, allv, side="right") / b.size).max())


def radial_profile(a, nbins=24):
    """Mean of the field in shells about the grid centre; returns (profile, cells per shell)."""
    L = a.shape[0]
    z, y, x = np.indices(a.shape)
    r = np.sqrt((x - L / 2 + 0.5) ** 2 + (y - L / 2 + 0.5) ** 2 + (z - L / 2 + 0.5) ** 2)
    idx = np.minimum((r / (L / 2) * nbins).astype(int), nbins - 1).ravel()
    cnt = np.bincount(idx, minlength=nbins)
    return np.bincount(idx, a.ravel().astype(np.float64), nbins) / np.maximum(cnt, 1), cnt


def settings(sim, L, prec, noise=0.0, phys=PHYS):
    return sim.Settings(L=L, precision=prec, backend="B200", kernel_language="Plain", noise=noise, **phys)


def make(sim, L, prec, noise=0.0, phys=PHYS, **kw):
    s = settings(sim, L, prec, noise, phys)
    mcd = sim.init_domain(s, 0, 1)
    return s, mcd, sim.init_fields(s, mcd, prec, **kw)


# ---------------------------------------------------------------- known answers
KAT = {
    ("Float32", 1): ("4434823e", "e2b4a83e"), ("Float32", 2): ("b9c7813e", "497fa83e"),
    ("Float64", 1): ("d8648d7a8846d03f", "2d95b7239c16d53f"), ("Float64", 2): ("32601913f738d03f", "206d7af5e80fd53f"),
}


@pytest.mark.parametrize("prec", ["Float32", "Float64"])
def test_kat_unit_simulation_iterate(sim, prec):
    """test/unit/simulation/unit-Simulation.jl:19-25 (L = 2, Settings() defaults) with the known
    answers of SURVEY.md Appendix C; step 2 also pins the alternating u ghost (Q1)."""
    s, mcd, f = make(sim, 2, prec, phys=PHYS_DEFAULT)
    for step in (1, 2):
        sim.iterate(f, s, mcd)
        u, v = sim.get_fields(f)
        assert np.unique(u).size == 1 and np.unique(v).size == 1
        assert u.tobytes()[:u.itemsize].hex() == KAT[(prec, step)][0]
        assert v.tobytes()[:v.itemsize].hex() == KAT[(prec, step)][1]
    f.close()


@pytest.mark.parametrize("prec", ["Float32", "Float64"])
@pytest.mark.parametrize("L", [8, 16, 32, 64, 128])
def test_init_fields_equals_cpu(sim, orc, prec, L):
    """unit-Simulation_CUDA.jl:10-32: fields.u ~ Array(fields_cuda.u) for L = 8..128 -- here exactly,
    ghosts and temporaries included."""
    s, mcd, f = make(sim, L, prec)
    o = orc.Oracle(L, prec, **PHYS)
    for slot in range(4):
        assert np.array_equal(np.asarray([f.u, f.v, f.u_temp, f.v_temp][slot]), o.get_ghosted(0, slot))
    assert f.u.shape == (L + 2, L + 2, L + 2)
    f.close()


# ---------------------------------------------------------------- noise = 0 trajectories
@pytest.mark.parametrize("prec", ["Float32", "Float64"])
def test_steps_bit_identical_to_oracle_L64(sim, orc, prec):
    """BASELINE config 1 physics, L = 64: after N = 1, 10, 100, 1000 steps the reference-arithmetic
    kernels equal the oracle bit for bit (so also within the 1e-5 / 1e-12 bars)."""
    s, mcd, f = make(sim, 64, prec)
    o = orc.Oracle(64, prec, **PHYS)
    done = 0
    for n in (1, 10, 100, 1000):
        sim.iterate(f, s, mcd, nsteps=n - done)
        o.step(n - done)
        done = n
        u, v = sim.get_fields(f)
        uo, vo = o.get_fields()
        assert rel_err(u, uo) <= REL_TOL[prec] and rel_err(v, vo) <= REL_TOL[prec]
        assert np.array_equal(u, uo) and np.array_equal(v, vo), f"{prec}: not bit-identical after {n} steps"
    f.close()


def test_native_f32_within_tolerance_L64_1000(sim, orc):
    """Pure-Float32 FMA arithmetic against the (mixed-precision) reference: the 1e-5 bar after 1000 steps."""
    s, mcd, f = make(sim, 64, "Float32", arith=NATIVE)
    o = orc.Oracle(64, "Float32", **PHYS)
    done = 0
    for n in (1, 10, 100, 1000):
        sim.iterate(f, s, mcd, nsteps=n - done)
        o.step(n - done)
        done = n
        u, v = sim.get_fields(f)
        uo, vo = o.get_fields()
        assert rel_err(u, uo) <= 1e-5 and rel_err(v, vo) <= 1e-5, (n, rel_err(u, uo), rel_err(v, vo))
    f.close()


@pytest.mark.parametrize("prec,arith", [("Float32", REFERENCE), ("Float32", NATIVE), ("Float64", REFERENCE)])
def test_config2_L256(sim, orc, prec, arith):
    """BASELINE config 2: L = 256, noise = 0, one B200 vs the reference CPU solver, N = 1, 10, 100, 1000."""
    s, mcd, f = make(sim, 256, prec, arith=arith)
    o = orc.Oracle(256, prec, **PHYS)
    done = 0
    for n in (1, 10, 100, 1000):
        sim.iterate(f, s, mcd, nsteps=n - done)
        o.step(n - done)
        done = n
        u, v = sim.get_fields(f)
        uo, vo = o.get_fields()
        assert rel_err(u, uo) <= REL_TOL[prec] and rel_err(v, vo) <= REL_TOL[prec]
        if arith == REFERENCE:
            assert np.array_equal(u, uo) and np.array_equal(v, vo)
    f.close()


# ---------------------------------------------------------------- every cell, ragged sizes
def custom_fields(sim, sizes, prec, noise=0.0, offsets=(0, 0, 0), arith=REFERENCE, seed=77, phys=PHYS):
    """A handle for an arbitrary (sx, sy, sz) subdomain sitting at `offsets` of a larger even global box."""
    s = sim.Settings(L=2, precision=prec, backend="B200", noise=noise, **phys)
    glob = tuple(int(2 * ((o + n + 2) // 2 + 1)) for o, n in zip(offsets, sizes))
    mcd = sim.MPICartDomain(rank=0, nranks=1, dims=(1, 1, 1), coords=(0, 0, 0), proc_sizes=tuple(sizes),
                            proc_offsets=tuple(offsets))
    return s, mcd, sim.init_fields(s, mcd, prec, arith=arith, seed=seed, global_extent=glob)


@pytest.mark.parametrize("prec", ["Float32", "Float64"])
@pytest.mark.parametrize("sizes", [(1, 1, 1), (3, 2, 5), (37, 5, 9), (130, 17, 3), (128, 16, 64), (261, 33, 70)])
@pytest.mark.parametrize("kern", KERNELS)
def test_one_step_random_fields_all_kernels(sim, orc, prec, sizes, kern):
    """One calculate! on random u ~ U[0,1), v ~ U[0,0.5) (ghosts included): every cell carries
    non-trivial data; ragged sizes exercise tile edges, masks and TMA out-of-bounds fill."""
    T = np.float32 if prec == "Float32" else np.float64
    rng = np.random.default_rng(12345)
    sx, sy, sz = sizes
    u = rng.random((sz + 2, sy + 2, sx + 2)).astype(T)
    v = (rng.random((sz + 2, sy + 2, sx + 2)) * 0.5).astype(T)
    s, mcd, f = custom_fields(sim, sizes, prec)
    f.set_kernel(*kern)
    f.u.set(u)
    f.v.set(v)
    poison = np.full_like(u, 7.0)
    f.u_temp.set(poison)
    f.v_temp.set(poison)
    sim.calculate(f, s, mcd)
    ut, vt = np.asarray(f.u_temp), np.asarray(f.v_temp)
    eu, ev = orc.calculate_arrays(u, v, **PHYS)
    inner = (slice(1, -1),) * 3
    assert np.array_equal(ut[inner], eu[inner]) and np.array_equal(vt[inner], ev[inner])
    # calculate! writes interior cells only: the ghost layer keeps what it had
    ut[inner] = 7.0
    vt[inner] = 7.0
    assert (ut == 7.0).all() and (vt == 7.0).all()
    # inputs untouched
    assert np.array_equal(np.asarray(f.u), u) and np.array_equal(np.asarray(f.v), v)
    f.close()


@pytest.mark.parametrize("prec", ["Float32", "Float64"])
def test_kernel_variants_agree_bitwise(sim, prec):
    ref = None
    for kern in KERNELS:
        s, mcd, f = make(sim, 96, prec, noise=0.1)
        f.set_kernel(*kern)
        sim.iterate(f, s, mcd, nsteps=7)
        out = sim.get_fields(f)
        f.close()
        if ref is None:
            ref = out
        else:
            assert np.array_equal(out[0], ref[0]) and np.array_equal(out[1], ref[1]), kern


# ---------------------------------------------------------------- noise
@pytest.mark.parametrize("prec", ["Float32", "Float64"])
@pytest.mark.parametrize("offsets", [(0, 0, 0), (5, 11, 3)])
def test_noise_bit_identical_to_oracle_philox(sim, orc, prec, offsets):
    """With the same counter-based generator on the CPU, noisy steps are bit-comparable too."""
    T = np.float32 if prec == "Float32" else np.float64
    rng = np.random.default_rng(5)
    sizes = (45, 9, 6)
    shape = (sizes[2] + 2, sizes[1] + 2, sizes[0] + 2)
    u = rng.random(shape).astype(T)
    v = (rng.random(shape) * 0.5).astype(T)
    s, mcd, f = custom_fields(sim, sizes, prec, noise=0.1, offsets=offsets, seed=1234)
    f.u.set(u)
    f.v.set(v)
    f.step = 17
    sim.calculate(f, s, mcd)
    eu, ev = orc.calculate_arrays(u, v, noise=0.1, noise_mode=orc.NOISE_PHILOX, seed=1234, step=17, offsets=offsets, **PHYS)
    inner = (slice(1, -1),) * 3
    assert np.array_equal(np.asarray(f.u_temp)[inner], eu[inner])
    assert np.array_equal(np.asarray(f.v_temp)[inner], ev[inner])
    f.close()


@pytest.mark.parametrize("prec", ["Float32", "Float64"])
def test_noise_run_config1(sim, orc, prec):
    """BASELINE config 1 (examples/settings-files.toml: L = 64, noise = 0.1): 200 steps.
    (a) bit-identical to the oracle driven by the same Philox stream;
    (b) statistically equal to the oracle driven by xoshiro256++ streams like the reference's
        task-local RNG: mean, variance, extrema and pattern area of u and v."""
    s, mcd, f = make(sim, 64, prec, noise=0.1, seed=0x5EED)
    a = orc.Oracle(64, prec, noise=0.1, noise_mode=orc.NOISE_PHILOX, seed=0x5EED, **PHYS)
    b = orc.Oracle(64, prec, noise=0.1, noise_mode=orc.NOISE_XOSHIRO, seed=42, **PHYS)
    sim.iterate(f, s, mcd, nsteps=200)
    a.step(200)
    b.step(200)
    u, v = sim.get_fields(f)
    assert np.array_equal(u, a.get_fields()[0]) and np.array_equal(v, a.get_fields()[1])
    ub, vb = b.get_fields()
    for x, y in ((u, ub), (v, vb)):
        x, y = x.astype(np.float64), y.astype(np.float64)
        assert abs(x.mean() - y.mean()) < 2e-3
        assert abs(x.var() - y.var()) < 0.05 * max(y.var(), 1e-4)
        assert abs(x.min() - y.min()) < 0.1 and abs(x.max() - y.max()) < 0.1
    area, area_b = (v > 0.1).mean(), (vb > 0.1).mean()
    assert abs(area - area_b) < 0.02
    # distribution of values (Kolmogorov-Smirnov distance) and radial profile about the centre;
    # two independent noise realisations of the oracle differ by <= 0.005 / 0.003 / 0.035 here
    for x, y in ((u, ub), (v, vb)):
        assert ks_distance(x, y) < 0.02
        pa, cnt = radial_profile(x)
        pb, _ = radial_profile(y)
        assert (np.abs(pa - pb) * cnt).sum() / cnt.sum() < 0.01
        assert np.abs(pa - pb)[4:].max() < 0.08
    # on-device statistics agree with numpy
    st = f.stats(1, 0.1)
    assert st["count_above"] == float((v > 0.1).sum())
    assert abs(st["mean"] - v.astype(np.float64).mean()) < 1e-12 + 1e-9 * abs(st["mean"])
    assert st["min"] == float(v.min()) and st["max"] == float(v.max())
    f.close()


def test_native_noise_statistics(sim, orc):
    s, mcd, f = make(sim, 64, "Float32", noise=0.1, arith=NATIVE)
    o = orc.Oracle(64, "Float32", noise=0.1, noise_mode=orc.NOISE_PHILOX, seed=sim.DEFAULT_SEED, **PHYS)
    sim.iterate(f, s, mcd, nsteps=50)
    o.step(50)
    u, v = sim.get_fields(f)
    uo, vo = o.get_fields()
    # same random stream (to 2^-24), so the trajectories stay close
    assert rel_err(u, uo) < 1e-4 and rel_err(v, vo) < 1e-4
    f.close()


# ---------------------------------------------------------------- quirks, pieces, errors
def test_q1_global_boundary_ghosts_never_written(sim):
    s, mcd, f = make(sim, 16, "Float32")
    for step in range(1, 5):
        sim.iterate(f, s, mcd)
        g = np.asarray(f.u)
        want = 0.0 if step % 2 else 1.0
        assert (g[0] == want).all() and (g[-1] == want).all() and (g[:, 0] == want).all() and (g[:, :, -1] == want).all()
        assert (np.asarray(f.v)[0] == 0).all()
    f.close()


def test_exchange_calculate_swap_equals_iterate(sim):
    s, mcd, a = make(sim, 32, "Float64")
    _, _, b = make(sim, 32, "Float64")
    for _ in range(5):
        sim.exchange(a, mcd)
        sim.calculate(a, s, mcd)
        sim.swap(a)
    sim.iterate(b, s, mcd, nsteps=5)
    assert a.step == b.step == 5
    for x, y in zip(sim.get_fields(a), sim.get_fields(b)):
        assert np.array_equal(x, y)
    a.close()
    b.close()


def test_async_output_equals_blocking_get_fields(sim):
    """gsb_get_fields_async snapshots in stream order: steps enqueued after it must not leak into it."""
    import torch
    s, mcd, f = make(sim, 96, "Float32", noise=0.1)
    sim.iterate(f, s, mcd, nsteps=5)
    want_u, want_v = sim.get_fields(f)
    hu = torch.empty(96 ** 3, dtype=torch.float32).pin_memory()
    hv = torch.empty(96 ** 3, dtype=torch.float32).pin_memory()
    for _ in range(2):                                   # second round reuses the staging buffers
        sim.get_fields_async(f, hu.data_ptr(), hv.data_ptr())
        sim.iterate(f, s, mcd, nsteps=7)                 # keeps running while the copy drains
        sim.output_wait(f)
        assert np.array_equal(hu.numpy().reshape(96, 96, 96), want_u)
        assert np.array_equal(hv.numpy().reshape(96, 96, 96), want_v)
        want_u, want_v = sim.get_fields(f)
    f.close()


@pytest.mark.parametrize("prec", ["Float32", "Float64"])
def test_pdf_per_slice_histogram(sim, prec):
    """_compute_pdf of src/analysis/pdfcalc.jl:13-49 (as intended: the reference version is unfinished)
    against a numpy restatement evaluated in T."""
    T = np.float32 if prec == "Float32" else np.float64
    s, mcd, f = make(sim, 48, prec, noise=0.1)
    sim.iterate(f, s, mcd, nsteps=60)
    u, v = sim.get_fields(f)
    for slot, a, (lo, hi), nbins in ((0, u, (0.0, 1.2), 100), (1, v, (0.0, 0.5), 1000), (1, v, (0.05, 0.2), 7)):
        pdf, bins = f.pdf(slot, nbins, lo, hi)
        lo_t, hi_t = T(lo), T(hi)
        w = (hi_t - lo_t) / T(nbins)
        want = np.zeros((48, nbins), np.uint32)
        for z in range(48):
            x = a[z].ravel()
            x = x[(x >= lo_t) & (x <= hi_t)]
            b = np.floor((x - lo_t) / w).astype(np.int64)
            b[b >= nbins] = nbins - 1
            want[z] = np.bincount(b, minlength=nbins)
        assert np.array_equal(pdf, want)
        assert np.array_equal(bins, (lo_t + np.arange(nbins).astype(T) * w).astype(np.float64))
    one, _ = f.pdf(0, 1, 0.0, 1.0)
    assert (one == 48 * 48).all()                       # special case of the reference: slice_size
    f.close()


def test_errors(sim):
    with pytest.raises(sim.GsbError, match="odd"):
        make(sim, 7, "Float32")                         # Int64(L/2 - 6): InexactError in the reference
    with pytest.raises(sim.GsbError, match="Float32 only"):
        make(sim, 8, "Float64", arith=NATIVE)
    s, mcd, f = make(sim, 8, "Float32")
    with pytest.raises(sim.GsbError, match="not instantiated"):
        f.set_kernel(0, 5, 4, 64)
    with pytest.raises(sim.GsbError, match="not supported"):
        sim.iterate(f, sim.Settings(backend="CPU"), mcd)
    f.close()


@pytest.mark.parametrize("prec", ["Float32", "Float64"])
def test_calculate_arrays_on_caller_device_memory(sim, orc, prec):
    """The stateless entry on reference-layout device arrays (torch tensors stand in for CuArrays)."""
    import torch
    T = np.float32 if prec == "Float32" else np.float64
    rng = np.random.default_rng(3)
    sizes = (41, 13, 7)
    shape = (sizes[2] + 2, sizes[1] + 2, sizes[0] + 2)
    u = rng.random(shape).astype(T)
    v = (rng.random(shape) * 0.5).astype(T)
    du, dv = torch.from_numpy(u).cuda(), torch.from_numpy(v).cuda()
    ou, ov = torch.zeros_like(du), torch.zeros_like(dv)
    s = sim.Settings(noise=0.1, **PHYS)
    torch.cuda.synchronize()
    sim.calculate_arrays(du.data_ptr(), dv.data_ptr(), ou.data_ptr(), ov.data_ptr(), sizes, prec, s,
                         offsets=(3, 0, 9), seed=11, step=2)
    torch.cuda.synchronize()
    eu, ev = orc.calculate_arrays(u, v, noise=0.1, noise_mode=orc.NOISE_PHILOX, seed=11, step=2, offsets=(3, 0, 9), **PHYS)
    assert np.array_equal(ou.cpu().numpy(), eu) and np.array_equal(ov.cpu().numpy(), ev)


# ---------------------------------------------------------------- halo exchange (in-process ranks)
# halo schemes: the default pipelined one (update in z chunks, each chunk's new boundary values sent
# one step ahead) with several chunk counts, and the interior/shell split
SCHEMES = [("pipeline", None), ("pipeline", 1), ("pipeline", 2), ("pipeline", 3), ("pipeline", 5), ("split", None)]


@pytest.mark.parametrize("prec", ["Float32", "Float64"])
@pytest.mark.parametrize("nranks", [2, 4, 8])
@pytest.mark.parametrize("noise", [0.0, 0.1])
@pytest.mark.parametrize("scheme", SCHEMES)
def test_halo_exchange_bit_exact_vs_single_subdomain(sim, orc, monkeypatch, prec, nranks, noise, scheme):
    """N subdomains (the reference's Cartesian decomposition: 2 -> (2,1,1), 4 -> (2,2,1), 8 -> (2,2,2))
    exchanging ghost faces every step == one subdomain of the same global grid, bit for bit, and ==
    the oracle's N-rank run.  The ranks share cuda:0 here; the transport differs from the NCCL one
    only in who moves the packed faces."""
    monkeypatch.setenv("GSB_HALO", scheme[0])
    if scheme[1]:
        monkeypatch.setenv("GSB_HALO_CHUNKS", str(scheme[1]))
    else:
        monkeypatch.delenv("GSB_HALO_CHUNKS", raising=False)
    L, steps = 64, 25
    s = settings(sim, L, prec, noise)
    ranks = []
    for r in range(nranks):
        mcd = sim.init_domain(s, r, nranks)
        ranks.append((mcd, sim.init_fields(s, mcd, prec, seed=9)))
    sim.connect_local([f for _, f in ranks])
    sim.iterate_group([f for _, f in ranks], s, nsteps=steps)
    U = np.empty((L, L, L), ranks[0][1].dtype)
    V = np.empty_like(U)
    for mcd, f in ranks:
        u, v = sim.get_fields(f)
        ox, oy, oz = mcd.proc_offsets
        sx, sy, sz = mcd.proc_sizes
        U[oz:oz + sz, oy:oy + sy, ox:ox + sx] = u
        V[oz:oz + sz, oy:oy + sy, ox:ox + sx] = v
    _, mcd1, one = make(sim, L, prec, noise, seed=9)
    sim.iterate(one, s, mcd1, nsteps=steps)
    u1, v1 = sim.get_fields(one)
    assert np.array_equal(U, u1) and np.array_equal(V, v1)
    o = orc.Oracle(L, prec, dims=orc.dims_create(nranks), noise=noise, seed=9, **PHYS)
    o.step(steps)
    uo, vo = o.gather()
    assert np.array_equal(U, uo) and np.array_equal(V, vo)
    for _, f in ranks:
        f.close()
    one.close()


@pytest.mark.parametrize("prec", ["Float32", "Float64"])
@pytest.mark.parametrize("nranks,L", [(2, 2), (8, 2), (8, 4), (4, 6), (27, 6)])
@pytest.mark.parametrize("scheme", ["pipeline", "split"])
def test_halo_exchange_tiny_subdomains(sim, orc, monkeypatch, prec, nranks, L, scheme):
    """Subdomains of one to a few cells per side (every cell is a boundary cell; with L = 2 on 8 ranks
    each rank owns a single cell whose six neighbours are all ghosts)."""
    monkeypatch.setenv("GSB_HALO", scheme)
    monkeypatch.delenv("GSB_HALO_CHUNKS", raising=False)
    monkeypatch.delenv("GSB_KERNEL", raising=False)
    s = settings(sim, L, prec, 0.1, PHYS_DEFAULT)
    ranks = []
    for r in range(nranks):
        mcd = sim.init_domain(s, r, nranks)
        ranks.append((mcd, sim.init_fields(s, mcd, prec, seed=2)))
    sim.connect_local([f for _, f in ranks])
    sim.iterate_group([f for _, f in ranks], s, nsteps=5)
    o = orc.Oracle(L, prec, dims=orc.dims_create(nranks), noise=0.1, seed=2, **PHYS_DEFAULT)
    o.step(5)
    for r, (mcd, f) in enumerate(ranks):
        u, v = sim.get_fields(f)
        uo, vo = o.get_fields(r)
        assert np.array_equal(u, uo) and np.array_equal(v, vo), r
        f.close()


@pytest.mark.parametrize("nranks,L", [(3, 48), (6, 48), (12, 48), (27, 54)])
@pytest.mark.parametrize("scheme", [("pipeline", None), ("pipeline", 3), ("split", None)])
def test_halo_exchange_other_decompositions(sim, orc, monkeypatch, nranks, L, scheme):
    """Decompositions with interior ranks (two neighbours per dimension) and unequal dims:
    3 -> (3,1,1), 6 -> (3,2,1), 12 -> (3,2,2), 27 -> (3,3,3); against the oracle's N-rank run."""
    monkeypatch.setenv("GSB_HALO", scheme[0])
    monkeypatch.setenv("GSB_KERNEL", "tma:8:4:16")      # small subdomains default to the L1 kernel: keep the TMA one covered
    if scheme[1]:
        monkeypatch.setenv("GSB_HALO_CHUNKS", str(scheme[1]))
    else:
        monkeypatch.delenv("GSB_HALO_CHUNKS", raising=False)
    steps, prec = 9, "Float32"
    s = settings(sim, L, prec, 0.1)
    ranks = []
    for r in range(nranks):
        mcd = sim.init_domain(s, r, nranks)
        ranks.append((mcd, sim.init_fields(s, mcd, prec, seed=4)))
    sim.connect_local([f for _, f in ranks])
    sim.iterate_group([f for _, f in ranks], s, nsteps=steps)
    o = orc.Oracle(L, prec, dims=orc.dims_create(nranks), noise=0.1, seed=4, **PHYS)
    o.step(steps)
    for r, (mcd, f) in enumerate(ranks):
        u, v = sim.get_fields(f)
        uo, vo = o.get_fields(r)
        assert np.array_equal(u, uo) and np.array_equal(v, vo), r
        f.close()


@pytest.mark.parametrize("prec", ["Float32", "Float64"])
@pytest.mark.parametrize("nranks,L", [(2, 256), (4, 256), (8, 256), (3, 384)])
@pytest.mark.parametrize("chunks", [None, 1, 3])
@pytest.mark.parametrize("fuse", [2, 1, 0])
def test_x_faces_carried_by_the_update_kernel(sim, orc, monkeypatch, prec, nranks, L, chunks, fuse):
    """sx = 128 is a whole number of tiles, so the update kernel carries the x faces itself (default;
    GSB_FUSE_X=0 switches back to the strided pack / unpack kernels): the producer warps of the first / last
    x tile fetch the neighbour's column from the dense receive buffer and send the new boundary column with
    bulk copies.  With 3 ranks along x (Float32: one tile per rank) the middle rank's CTAs own both sides.
    Must equal the oracle's N-rank run bit for bit, and Array(fields.u) must still show the neighbour's
    column in the ghost layer."""
    monkeypatch.setenv("GSB_HALO", "pipeline")
    monkeypatch.setenv("GSB_FUSE_X", str(fuse))
    monkeypatch.setenv("GSB_KERNEL", "tma:8:4:16")      # the fused variant exists for the TMA kernel only
    if chunks:
        monkeypatch.setenv("GSB_HALO_CHUNKS", str(chunks))
    else:
        monkeypatch.delenv("GSB_HALO_CHUNKS", raising=False)
    steps = 7
    s = settings(sim, L, prec, 0.1)
    ranks = []
    for r in range(nranks):
        mcd = sim.init_domain(s, r, nranks)
        ranks.append((mcd, sim.init_fields(s, mcd, prec, seed=13)))
    assert ranks[0][0].proc_sizes[0] == 128
    sim.connect_local([f for _, f in ranks])
    launches0 = ranks[0][1].launch_count()
    sim.iterate_group([f for _, f in ranks], s, nsteps=3)
    sim.iterate_group([f for _, f in ranks], s, nsteps=steps - 3)
    per_step = (ranks[0][1].launch_count() - launches0) / steps
    o = orc.Oracle(L, prec, dims=orc.dims_create(nranks), noise=0.1, seed=13, **PHYS)
    o.step(steps)
    got = []
    for r, (mcd, f) in enumerate(ranks):
        u, v = sim.get_fields(f)
        uo, vo = o.get_fields(r)
        assert np.array_equal(u, uo) and np.array_equal(v, vo), r
        got.append(u)
    # rank 0's east ghost column is its east neighbour's column x = 0 (current values)
    mcd0, f0 = ranks[0]
    east = mcd0.proc_neighbors["east"]
    g = np.asarray(f0.u)
    assert np.array_equal(g[1:-1, 1:-1, -1], got[east][:, :, 0])
    if nranks in (2, 3) and fuse:
        # only x faces exist: no pack / unpack kernel at all after the first (priming) exchange
        C_ = chunks or (4 if ranks[0][0].proc_sizes[2] >= 192 else 2)
        assert per_step <= (1 if fuse == 1 else 3) * C_ + 2 * C_ / steps + 1e-9, per_step
    for _, f in ranks:
        f.close()


@pytest.mark.parametrize("scheme", ["pipeline", "split"])
def test_overlap_large_subdomains(sim, monkeypatch, scheme):
    """Subdomains big enough that the interior/shell split is active in every direction and that the
    pipelined scheme runs its default number of chunks; mixes piecewise calls and iterate!."""
    monkeypatch.setenv("GSB_HALO", scheme)
    monkeypatch.delenv("GSB_HALO_CHUNKS", raising=False)
    L, steps, nranks = 320, 6, 8
    s = settings(sim, L, "Float32", 0.1)
    ranks = []
    for r in range(nranks):
        mcd = sim.init_domain(s, r, nranks)
        ranks.append((mcd, sim.init_fields(s, mcd, "Float32", seed=9)))
    sim.connect_local([f for _, f in ranks])
    ranks[0][1].timers_enable(True)
    sim.iterate_group([f for _, f in ranks], s, nsteps=steps)
    t = ranks[0][1].timers()
    assert t["interior_ms"] > 0 and ranks[0][1].halo_bytes() == 3 * 2 * 160 * 160 * 4
    if scheme == "split":
        assert t["shell_ms"] > 0
    _, mcd1, one = make(sim, L, "Float32", 0.1, seed=9)
    sim.iterate(one, s, mcd1, nsteps=steps)
    u1, v1 = sim.get_fields(one)
    for mcd, f in ranks:
        u, v = sim.get_fields(f)
        ox, oy, oz = mcd.proc_offsets
        sx, sy, sz = mcd.proc_sizes
        assert np.array_equal(u, u1[oz:oz + sz, oy:oy + sy, ox:ox + sx])
        assert np.array_equal(v, v1[oz:oz + sz, oy:oy + sy, ox:ox + sx])
        f.close()
    one.close()


# ---------------------------------------------------------------- full size
def test_full_size_L1024_matches_small_grid_in_the_domain_of_dependence(sim, orc):
    """BASELINE config 3 size (L = 1024, Float32).  After n steps nothing farther than n cells from the
    initial block or from the boundary has changed, so (i) the block's neighbourhood must equal, bit
    for bit, the same neighbourhood of a small grid the oracle can compute, and (ii) everything else
    away from the boundary is still exactly u = 1, v = 0."""
    L, n, small = 1024, 12, 64
    s, mcd, f = make(sim, L, "Float32")
    sim.iterate(f, s, mcd, nsteps=n)
    u, v = sim.get_fields(f)
    f.close()
    o = orc.Oracle(small, "Float32", **PHYS)
    o.step(n)
    uo, vo = o.get_fields()
    r = 6 + n + 1
    big = slice(L // 2 - r, L // 2 + r + 1)
    sm = slice(small // 2 - r, small // 2 + r + 1)
    assert np.array_equal(u[big, big, big], uo[sm, sm, sm]) and np.array_equal(v[big, big, big], vo[sm, sm, sm])
    far = np.ones_like(u, dtype=bool)
    far[big, big, big] = False
    far[:n + 1] = far[-n - 1:] = False
    far[:, :n + 1] = far[:, -n - 1:] = False
    far[:, :, :n + 1] = far[:, :, -n - 1:] = False
    assert (u[far] == 1).all() and (v[far] == 0).all()


def test_full_size_L1024_noise_statistics(sim, orc):
    """BASELINE config 3 with noise = 0.1 (the headline bench workload), native Float32 arithmetic.  The
    fields are far too large to compare cell by cell with the CPU solver, but away from the tiny
    initial block the update is a linear stochastic recursion whose moments do not depend on L: mean and
    variance of u and v after n steps must match a small grid run by the oracle with xoshiro noise (the
    reference's kind of generator), to sampling accuracy; computed on the device (no 8 GB download)."""
    n = 6
    s, mcd, f = make(sim, 1024, "Float32", noise=0.1, arith=NATIVE)
    sim.iterate(f, s, mcd, nsteps=n)
    su, sv = f.stats(0, 0.5), f.stats(1, 0.1)
    f.close()
    o = orc.Oracle(128, "Float32", noise=0.1, noise_mode=orc.NOISE_XOSHIRO, seed=77, **PHYS)
    o.step(n)
    uo, vo = (a.astype(np.float64) for a in o.get_fields())
    # reference region: a corner block of the small grid that neither the boundary layer (8 cells) nor
    # the initial block's neighbourhood (n cells around the centre) has reached -- pure noise statistics
    far = (slice(8, 40),) * 3
    assert abs(su["mean"] - uo[far].mean()) < 2e-3
    assert abs(su["var"] - uo[far].var()) < 0.05 * uo[far].var()
    # v is non-zero only around the initial block: its total mass does not depend on L
    assert abs(sv["sum"] - vo.sum()) < 0.03 * vo.sum()
    assert su["min"] > -0.5 and su["max"] < 2.0 and sv["min"] >= -1e-3


def test_bench_own_arm_prints_one_json_line():
    """bench.py on a small grid: one JSON line with every key of the contract."""
    import json
    import os
    import subprocess
    import sys
    from conftest import ROOT
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--L", "128", "--steps", "20", "--warmup", "3",
                        "--e2e-steps", "2"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, r.stdout
    d = json.loads(lines[0])
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "clocks", "gpu_launches", "roofline", "e2e", "cpu_baseline"):
        assert key in d, key
    assert d["n_gpus"] == 1 and d["steps"] == 20 and d["gpu_launches"] == 20 and d["value"] > 1.0
    for key in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
        assert key in d["roofline"], key
    assert d["roofline"]["bound"] == "hbm" and abs(d["roofline"]["frac"] - d["roofline"]["achieved"] / d["roofline"]["peak"]) < 1e-9
    assert d["e2e"]["h2d_bytes_per_step"] > 0 and d["e2e"]["d2h_bytes_per_step"] > 0 and d["e2e"]["value"] > 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["value"] > 0
    # round 2: the K steps after warm-up are window 0 = value; later windows and the other arithmetic beside it
    w = d["windows"]
    assert w["n"] == 3 and w["value_window"] == 0 and abs(w["ms_per_step"][0] - d["ms_per_step"]) < 1e-12
    assert len(w["ms_per_step_by_rank_all"]) == 3 and len(w["ms_per_step_by_rank"]) == 1
    assert d["value_reference_arith"]["value"] > 0 and d["e2e_async"]["value"] > 0 and d["e2e_every_step"]["value"] > 0
    assert d["config"]["scaling_mode"] == "weak" and d["config"]["per_gpu"] == [128, 128, 128]


def test_bench_strong_scaling_mode_one_gpu():
    """--scaling strong --global-L: sizes = L ./ dims (communication.jl:73); at one GPU the whole cube."""
    import json
    import os
    import subprocess
    import sys
    from conftest import ROOT
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--scaling", "strong", "--global-L", "96", "--precision",
                        "Float64", "--steps", "5", "--warmup", "3", "--windows", "1", "--no-e2e", "--no-cpu-baseline"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads([ln for ln in r.stdout.splitlines() if ln.strip()][0])
    assert d["scaling"] == "strong" and d["dtype"] == "f64" and d["config"]["global"] == [96, 96, 96]
    assert d["config"]["per_gpu"] == [96, 96, 96] and d["value"] > 0 and "value_reference_arith" not in d
