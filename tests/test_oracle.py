"""CPU tests that pin the oracle (oracle/gs_oracle.c).

The reference cannot run here (Julia) and its tests hold no golden vector for calculate!/exchange!,
so the oracle is pinned by: the hand-derived known answers for the reference's own unit-test case
(SURVEY.md Appendix C), an independent numpy-scalar restatement of the Julia expression, and the
invariants of init_fields / the decomposition.
"""
import numpy as np
import pytest

from conftest import PHYS, PHYS_DEFAULT


# ---------------------------------------------------------------- known answers (Appendix C)
# Case: test/unit/simulation/unit-Simulation.jl:19-25 -- Settings() defaults, L = 2, 1 rank.
KAT = {
    ("Float32", 1): ("4434823e", "e2b4a83e"),
    ("Float32", 2): ("b9c7813e", "497fa83e"),
    ("Float64", 1): ("d8648d7a8846d03f", "2d95b7239c16d53f"),
    ("Float64", 2): ("32601913f738d03f", "206d7af5e80fd53f"),
}


@pytest.mark.parametrize("prec", ["Float32", "Float64"])
def test_kat_unit_simulation_iterate(orc, prec):
    o = orc.Oracle(2, prec, **PHYS_DEFAULT)
    for step in (1, 2):
        o.step(1)
        u, v = o.get_fields()
        assert np.unique(u).size == 1 and np.unique(v).size == 1     # all 8 cells identical
        isz = u.itemsize
        assert u.tobytes()[:isz].hex() == KAT[(prec, step)][0]
        assert v.tobytes()[:isz].hex() == KAT[(prec, step)][1]


# ---------------------------------------------------------------- independent restatement
def julia_cell(T, u, v, i, j, k, Du, Dv, F, K, dt, noise, r):
    """Common.jl:13-18 + Simulation_CPU.jl:96-109 with numpy scalars standing in for Julia's:
    T op T stays T, T op Float64 promotes to Float64.  Arrays indexed [k, j, i]."""
    f64 = np.float64

    def lap(a):
        s = a[k, j, i - 1] + a[k, j, i + 1]       # left fold, all in T
        s = s + a[k, j - 1, i]
        s = s + a[k, j + 1, i]
        s = s + a[k - 1, j, i]
        s = s + a[k + 1, j, i]
        ll = f64(s) - f64(6.0) * f64(a[k, j, i])    # `- 6.0 * var[i,j,k]` promotes
        return ll / f64(6.0)

    uc, vc = u[k, j, i], v[k, j, i]
    Du, Dv, F, K, dt, noise = (T(x) for x in (Du, Dv, F, K, dt, noise))
    uvv = uc * (vc * vc)                            # T
    du = f64(Du) * lap(u) - f64(uvv)
    du = du + f64(F) * (f64(1.0) - f64(uc))
    du = du + f64(noise) * f64(r)
    dv = f64(Dv) * lap(v) + f64(uvv)
    dv = dv - f64((F + K) * vc)                     # (F + K) * v in T
    return T(f64(uc) + du * f64(dt)), T(f64(vc) + dv * f64(dt))


@pytest.mark.parametrize("T", [np.float32, np.float64])
def test_oracle_matches_numpy_scalar_restatement(orc, T):
    rng = np.random.default_rng(12345)
    shape = (6, 7, 8)                               # ghosted [z, y, x]: interior 4 x 5 x 6
    u = rng.random(shape).astype(T)
    v = (rng.random(shape) * 0.5).astype(T)
    phys = dict(Du=0.2, Dv=0.1, F=0.02, k=0.048, dt=0.7)
    ut, vt = orc.calculate_arrays(u, v, **phys)
    with np.errstate(all="ignore"):
        for k in range(1, shape[0] - 1):
            for j in range(1, shape[1] - 1):
                for i in range(1, shape[2] - 1):
                    eu, ev = julia_cell(T, u, v, i, j, k, phys["Du"], phys["Dv"], phys["F"], phys["k"],
                                        phys["dt"], 0.0, 0.0)
                    assert ut[k, j, i].tobytes() == eu.tobytes(), (k, j, i)
                    assert vt[k, j, i].tobytes() == ev.tobytes(), (k, j, i)
    # calculate! writes interior cells only
    assert ut[0].max() == 0 and ut[:, 0].max() == 0 and ut[:, :, 0].max() == 0


@pytest.mark.parametrize("T", [np.float32, np.float64])
def test_oracle_philox_noise_matches_restatement(orc, T):
    rng = np.random.default_rng(7)
    shape = (5, 5, 9)
    u = rng.random(shape).astype(T)
    v = (rng.random(shape) * 0.5).astype(T)
    phys = dict(Du=0.2, Dv=0.1, F=0.02, k=0.048, dt=1.0)
    off = (5, 11, 3)                                 # unaligned global x offset
    ut, vt = orc.calculate_arrays(u, v, noise=0.1, noise_mode=orc.NOISE_PHILOX, seed=99, step=4, offsets=off, **phys)
    for k in range(1, shape[0] - 1):
        for j in range(1, shape[1] - 1):
            for i in range(1, shape[2] - 1):
                r = orc.noise_r(99, 4, off[0] + i - 1, off[1] + j - 1, off[2] + k - 1)
                eu, ev = julia_cell(T, u, v, i, j, k, phys["Du"], phys["Dv"], phys["F"], phys["k"], phys["dt"],
                                    0.1, r)
                assert ut[k, j, i].tobytes() == eu.tobytes()
                assert vt[k, j, i].tobytes() == ev.tobytes()


# ---------------------------------------------------------------- Philox
def test_philox_known_answers(orc):
    # Random123 kat_vectors, philox4x32 with 10 rounds
    assert orc.philox4x32_10((0, 0, 0, 0), (0, 0)) == (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)
    assert orc.philox4x32_10((0xffffffff,) * 4, (0xffffffff,) * 2) == (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)
    assert orc.philox4x32_10((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0)) == (
        0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)


def test_noise_variate_definition_and_moments(orc):
    # r = -1 + 2 * word * 2^-32 with word = philox(ctr=(gx>>2, gy, gz, step), key=seed)[gx & 3]
    seed = 0x0123456789ABCDEF
    for gx, gy, gz, st in [(0, 0, 0, 0), (5, 1, 2, 3), (1023, 77, 1000, 999)]:
        w = orc.philox4x32_10((gx >> 2, gy, gz, st), (seed & 0xffffffff, seed >> 32))[gx & 3]
        assert orc.noise_r(seed, st, gx, gy, gz) == -1.0 + 2.0 * (w * 2.0 ** -32)
    r = np.array([orc.noise_r(1, s, x, y, 3) for s in range(4) for y in range(50) for x in range(200)])
    assert r.min() >= -1.0 and r.max() < 1.0
    n = r.size
    assert abs(r.mean()) < 4 * np.sqrt(1 / 3 / n)              # mean 0
    assert abs(r.var() - 1 / 3) < 0.01                          # Uniform(-1,1): var 1/3
    assert np.unique(r).size == n                               # no (step, cell) collisions


# ---------------------------------------------------------------- init_fields
@pytest.mark.parametrize("L", [8, 16, 32, 64, 128])            # unit-Simulation_CUDA.jl:26-30
@pytest.mark.parametrize("prec", ["Float32", "Float64"])
def test_init_fields(orc, L, prec):
    o = orc.Oracle(L, prec, **PHYS)
    T = o.np_dtype
    u, v = o.get_ghosted(0, 0), o.get_ghosted(0, 1)
    lo, hi = max(L // 2 - 6, 0), min(L // 2 + 6, L - 1)
    n_block = (hi - lo + 1) ** 3
    assert n_block == (2197 if L >= 16 else L ** 3)
    assert (u == T(0.25)).sum() == n_block and (v == T(0.33)).sum() == n_block
    assert (u == T(1)).sum() == u.size - n_block and (v == T(0)).sum() == v.size - n_block
    blk = u[lo + 1:hi + 2, lo + 1:hi + 2, lo + 1:hi + 2]       # +1: ghost layer
    assert (blk == T(0.25)).all()
    assert (o.get_ghosted(0, 2) == 0).all() and (o.get_ghosted(0, 3) == 0).all()   # temps are zeros, ghosts too


# ---------------------------------------------------------------- quirks the backend must keep
def test_q1_alternating_u_ghost(orc):
    """u = ones / u_temp = zeros include the ghosts and nothing ever writes global-boundary ghosts,
    so after the swap u's ghost layer is 0 on odd step counts and 1 on even ones."""
    o = orc.Oracle(8, "Float32", **PHYS)
    for step in range(1, 5):
        o.step(1)
        g = o.get_ghosted(0, 0)
        assert (g[0] == (0 if step % 2 else 1)).all() and (g[:, :, -1] == (0 if step % 2 else 1)).all()
        assert (o.get_ghosted(0, 1)[0] == 0).all()


def test_noise_zero_draw_does_not_change_results(orc):
    """The reference draws a variate per cell even at noise = 0 (Simulation_CPU.jl:103); adding
    0.0 * r must not change a bit, so the oracle may skip the draw."""
    a = orc.Oracle(16, "Float32", noise=0.0, noise_mode=orc.NOISE_SKIP, **PHYS)
    b = orc.Oracle(16, "Float32", noise=0.0, noise_mode=orc.NOISE_XOSHIRO, **PHYS)
    a.step(20)
    b.step(20)
    for x, y in zip(a.get_fields(), b.get_fields()):
        assert np.array_equal(x, y)


def test_dims_create(orc):
    assert [orc.dims_create(n) for n in (1, 2, 4, 8)] == [(1, 1, 1), (2, 1, 1), (2, 2, 1), (2, 2, 2)]
    for n in (3, 6, 12, 16):
        d = orc.dims_create(n)
        assert d[0] * d[1] * d[2] == n and d[0] >= d[1] >= d[2]


def test_uneven_L_rejected(orc):
    with pytest.raises(ValueError):
        orc.Oracle(10, "Float32", dims=(4, 1, 1), **PHYS)     # communication.jl:73 -> InexactError


@pytest.mark.parametrize("prec", ["Float32", "Float64"])
@pytest.mark.parametrize("nranks", [2, 4, 8])
@pytest.mark.parametrize("noise", [0.0, 0.1])
def test_decomposition_invariance(orc, prec, nranks, noise):
    """N in-process ranks exchanging faces == 1 rank, bit for bit (Philox is keyed on the global cell)."""
    L, steps = 16, 12
    one = orc.Oracle(L, prec, noise=noise, seed=3, **PHYS)
    many = orc.Oracle(L, prec, dims=orc.dims_create(nranks), noise=noise, seed=3, **PHYS)
    assert many.nranks == nranks
    one.step(steps)
    many.step(steps)
    for a, b in zip(one.gather(), many.gather()):
        assert np.array_equal(a, b)


def test_rank_geometry_matches_reference_rules(orc):
    o = orc.Oracle(16, "Float32", dims=(2, 2, 2), **PHYS)
    info = o.rank_info(5)                                       # coords (1, 0, 1): rank = (cx*dy + cy)*dz + cz
    assert info["coords"] == (1, 0, 1) and info["sizes"] == (8, 8, 8) and info["offsets"] == (8, 0, 8)
    # west, east, down, up, south, north ; -1 = MPI.PROC_NULL (non-periodic)
    assert info["neighbors"] == (1, -1, -1, 7, 4, -1)


def test_axis_symmetry(orc):
    """The noise-free problem is symmetric under permutations of x, y, z up to rounding (the
    left-fold order differs per axis)."""
    o = orc.Oracle(24, "Float64", **PHYS)
    o.step(30)
    u, v = o.get_fields()
    for a in (u, v):
        assert np.allclose(a, a.transpose(2, 1, 0), rtol=0, atol=1e-13)
        assert np.allclose(a, a.transpose(0, 2, 1), rtol=0, atol=1e-13)
