"""gsb_field_stats_noalloc -- the statistics entry point for callers that may not allocate.

Collected last on purpose (file name), so a `-x` run has already been through every other test when
it gets here.  The caller's side is played by torch tensors: the field lives in a device array the
"caller" owns (reference layout, ghosted), the five results land in a caller-owned device buffer,
and the library itself allocates nothing -- its per-plane scratch is a fixed array in the library
image.  The check: same numbers as the handle-based gsb_field_stats, on two calls in a row, so that
whatever the first call left in the shared scratch must not leak into the second.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import PHYS

pytestmark = pytest.mark.gpu


def call_noalloc(pkg, dtype_code, sizes, field_t, thresh, out_t, stream=None):
    err = C.create_string_buffer(256)
    rc = pkg._lib.lib().gsb_field_stats_noalloc(dtype_code, (C.c_int64 * 3)(*sizes), field_t.data_ptr(), thresh,
                                                out_t.data_ptr(), stream, err, len(err))
    return rc, err.value.decode()


@pytest.mark.parametrize("prec", ["Float32", "Float64"])
def test_noalloc_stats_match_handle_stats_twice_in_a_row(pkg, sim, prec):
    import torch
    L = 48
    s = sim.Settings(L=L, precision=prec, backend="B200", noise=0.1, **PHYS)
    mcd = sim.init_domain(s, 0, 1)
    f = sim.init_fields(s, mcd, prec, seed=21)
    sim.iterate(f, s, mcd, nsteps=40)
    code = pkg._lib.GSB_F32 if prec == "Float32" else pkg._lib.GSB_F64
    # the caller's own device arrays (what a CuArray of Fields.u / Fields.v would be)
    u_dev = torch.from_numpy(np.asarray(f.u)).cuda()
    v_dev = torch.from_numpy(np.asarray(f.v)).cuda()
    out_u = torch.full((5,), -1.0, dtype=torch.float64, device="cuda")
    out_v = torch.full((5,), -1.0, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    st = torch.cuda.current_stream().cuda_stream
    # two calls back to back on one stream: u (large values everywhere) then v (mostly zeros)
    rc, msg = call_noalloc(pkg, code, (L, L, L), u_dev, 0.5, out_u, st)
    assert rc == 0, msg
    rc, msg = call_noalloc(pkg, code, (L, L, L), v_dev, 0.1, out_v, st)
    assert rc == 0, msg
    torch.cuda.synchronize()
    for slot, thresh, out in ((pkg._lib.GSB_U, 0.5, out_u), (pkg._lib.GSB_V, 0.1, out_v)):
        want = f.stats(slot, thresh)
        got = out.cpu().numpy()
        # tolerance of the statistics checks in test_gpu_parity.py: sums to 1e-9 relative
        # (the two entry points add the partial sums in a different order), the rest exact
        assert abs(got[0] - want["sum"]) <= 1e-12 + 1e-9 * abs(want["sum"])
        assert abs(got[1] - want["sumsq"]) <= 1e-12 + 1e-9 * abs(want["sumsq"])
        assert got[2] == want["min"] and got[3] == want["max"] and got[4] == want["count_above"]
    # and against numpy on the host copy
    v_host = np.asarray(f.v)[1:-1, 1:-1, 1:-1].astype(np.float64)
    got = out_v.cpu().numpy()
    assert abs(got[0] - v_host.sum()) <= 1e-9 * abs(v_host.sum()) + 1e-12
    assert got[4] == float((v_host > 0.1).sum())
    f.close()


def test_noalloc_rejects_fields_beyond_the_builtin_scratch(pkg):
    import torch
    dummy = torch.zeros(8, dtype=torch.float32, device="cuda")
    out = torch.zeros(5, dtype=torch.float64, device="cuda")
    too_many = 4097                                            # GSB_NOALLOC_MAX_PLANES + 1
    rc, msg = call_noalloc(pkg, pkg._lib.GSB_F32, (1, 1, too_many), dummy, 0.0, out)
    assert rc != 0 and "GSB_NOALLOC_MAX_PLANES" in msg and "4096" in msg
    rc, msg = call_noalloc(pkg, 7, (4, 4, 4), dummy, 0.0, out)
    assert rc != 0 and "dtype" in msg
    # like every other entry point, the reason is also what gsb_last_error() returns (a caller may pass no errbuf)
    assert pkg._lib.lib().gsb_last_error().decode() == msg


def test_noalloc_ragged_box(pkg):
    """Non-cubic, odd sizes; compares with numpy."""
    import torch
    rng = np.random.default_rng(8)
    sx, sy, sz = 37, 5, 9
    a = rng.standard_normal((sz + 2, sy + 2, sx + 2))
    dev = torch.from_numpy(a).cuda()
    out = torch.zeros(5, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    rc, msg = call_noalloc(pkg, pkg._lib.GSB_F64, (sx, sy, sz), dev, 0.25, out, torch.cuda.current_stream().cuda_stream)
    assert rc == 0, msg
    torch.cuda.synchronize()
    inner = a[1:-1, 1:-1, 1:-1]
    got = out.cpu().numpy()
    assert abs(got[0] - inner.sum()) < 1e-10 and abs(got[1] - (inner ** 2).sum()) < 1e-10
    assert got[2] == inner.min() and got[3] == inner.max() and got[4] == float((inner > 0.25).sum())
