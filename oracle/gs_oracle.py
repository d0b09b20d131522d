"""ctypes binding of the CPU oracle (oracle/gs_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libgs_oracle.so")

F32, F64 = 0, 1
NOISE_SKIP, NOISE_XOSHIRO, NOISE_PHILOX = 0, 1, 2
_NP = {F32: np.float32, F64: np.float64}


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "gs_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE, "-B", "libgs_oracle.so"], check=True, capture_output=True)
    return _LIB_PATH


_lib = None


def dtype_code(dtype) -> int:
    if isinstance(dtype, str):
        return {"float32": F32, "f32": F32, "float64": F64, "f64": F64}[dtype.lower()]
    if isinstance(dtype, int):
        return dtype
    return F32 if np.dtype(dtype) == np.float32 else F64


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = C.CDLL(_LIB_PATH)
        i64p, ip, vp = C.POINTER(C.c_int64), C.POINTER(C.c_int), C.c_void_p
        L.gso_create.restype = vp
        L.gso_create.argtypes = [C.c_int, C.c_int64, ip] + [C.c_double] * 6 + [C.c_int, C.c_uint64]
        L.gso_destroy.argtypes = [vp]
        L.gso_nranks.argtypes = [vp]
        L.gso_step_count.restype = C.c_int64
        L.gso_step_count.argtypes = [vp]
        L.gso_set_step_count.argtypes = [vp, C.c_int64]
        L.gso_rank_info.argtypes = [vp, C.c_int, i64p, i64p, ip, ip]
        L.gso_init_fields.argtypes = [vp]
        L.gso_calculate.argtypes = [vp]
        L.gso_exchange.argtypes = [vp]
        L.gso_swap.argtypes = [vp]
        L.gso_step.argtypes = [vp, C.c_int]
        L.gso_get_ghosted.argtypes = [vp, C.c_int, C.c_int, vp]
        L.gso_set_ghosted.argtypes = [vp, C.c_int, C.c_int, vp]
        L.gso_get_fields.argtypes = [vp, C.c_int, vp, vp]
        L.gso_gather.argtypes = [vp, vp, vp]
        L.gso_dims_create.argtypes = [C.c_int, ip]
        L.gso_calculate_arrays.argtypes = ([C.c_int, i64p, i64p, vp, vp, vp, vp] + [C.c_double] * 6 +
                                           [C.c_int, C.c_uint64, C.c_uint32])
        L.gso_philox4x32_10.argtypes = [C.POINTER(C.c_uint32)] * 3
        L.gso_noise_r.restype = C.c_double
        L.gso_noise_r.argtypes = [C.c_uint64] + [C.c_uint32] * 4
        L.gso_max_threads.restype = C.c_int
        L.gso_set_threads.argtypes = [C.c_int]
        _lib = L
    return _lib


def dims_create(n: int) -> tuple[int, int, int]:
    d = (C.c_int * 3)()
    lib().gso_dims_create(n, d)
    return tuple(d)


def philox4x32_10(ctr, key):
    c = (C.c_uint32 * 4)(*ctr)
    k = (C.c_uint32 * 2)(*key)
    o = (C.c_uint32 * 4)()
    lib().gso_philox4x32_10(c, k, o)
    return tuple(o)


def noise_r(seed, step, gx, gy, gz) -> float:
    return lib().gso_noise_r(seed, step, gx, gy, gz)


def calculate_arrays(u, v, *, Du, Dv, F, k, dt, noise=0.0, noise_mode=NOISE_SKIP, seed=0, step=0,
                     offsets=(0, 0, 0)):
    """One calculate! on ghosted numpy arrays indexed [z, y, x] (x fastest).  Returns (u_temp, v_temp)
    with ghosts zero (calculate! writes interior cells only)."""
    assert u.dtype == v.dtype and u.shape == v.shape and u.flags.c_contiguous and v.flags.c_contiguous
    dt_code = F32 if u.dtype == np.float32 else F64
    nz, ny, nx = u.shape
    size = (C.c_int64 * 3)(nx - 2, ny - 2, nz - 2)
    off = (C.c_int64 * 3)(*offsets)
    ut, vt = np.zeros_like(u), np.zeros_like(v)
    rc = lib().gso_calculate_arrays(dt_code, size, off, u.ctypes.data, v.ctypes.data, ut.ctypes.data,
                                    vt.ctypes.data, Du, Dv, F, k, dt, noise, noise_mode, seed, step)
    if rc:
        raise ValueError("noise_mode SKIP requires noise == 0")
    return ut, vt


class Oracle:
    """N in-process ranks of the reference CPU solver.  Arrays are numpy, indexed [z, y, x]."""

    def __init__(self, L, dtype="Float32", dims=(1, 1, 1), *, Du, Dv, F, k, dt, noise=0.0,
                 noise_mode=None, seed=0):
        self.dtype = dtype_code(dtype)
        self.np_dtype = _NP[self.dtype]
        if noise_mode is None:
            noise_mode = NOISE_SKIP if noise == 0.0 else NOISE_PHILOX
        d = (C.c_int * 3)(*dims)
        self.L, self.dims = L, tuple(dims)
        self._h = lib().gso_create(self.dtype, L, d, Du, Dv, F, k, dt, noise, noise_mode, seed)
        if not self._h:
            raise ValueError(f"L={L} must divide evenly by dims={dims} (reference: InexactError)")
        self.nranks = lib().gso_nranks(self._h)
        lib().gso_init_fields(self._h)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().gso_destroy(self._h)
            self._h = None

    def rank_info(self, rank):
        size, off = (C.c_int64 * 3)(), (C.c_int64 * 3)()
        coords, nbr = (C.c_int * 3)(), (C.c_int * 6)()
        lib().gso_rank_info(self._h, rank, size, off, coords, nbr)
        return dict(sizes=tuple(size), offsets=tuple(off), coords=tuple(coords), neighbors=tuple(nbr))

    @property
    def step_count(self):
        return lib().gso_step_count(self._h)

    def init_fields(self):
        lib().gso_init_fields(self._h)

    def exchange(self):
        lib().gso_exchange(self._h)

    def calculate(self):
        if lib().gso_calculate(self._h):
            raise ValueError("noise_mode SKIP requires noise == 0")

    def swap(self):
        lib().gso_swap(self._h)

    def step(self, n=1):
        if lib().gso_step(self._h, n):
            raise ValueError("noise_mode SKIP requires noise == 0")

    def _shape(self, rank):
        sx, sy, sz = self.rank_info(rank)["sizes"]
        return sx, sy, sz

    def get_ghosted(self, rank=0, which=0):
        sx, sy, sz = self._shape(rank)
        a = np.empty((sz + 2, sy + 2, sx + 2), self.np_dtype)
        lib().gso_get_ghosted(self._h, rank, which, a.ctypes.data)
        return a

    def set_ghosted(self, a, rank=0, which=0):
        a = np.ascontiguousarray(a, self.np_dtype)
        lib().gso_set_ghosted(self._h, rank, which, a.ctypes.data)

    def get_fields(self, rank=0):
        sx, sy, sz = self._shape(rank)
        u = np.empty((sz, sy, sx), self.np_dtype)
        v = np.empty_like(u)
        lib().gso_get_fields(self._h, rank, u.ctypes.data, v.ctypes.data)
        return u, v

    def gather(self):
        u = np.empty((self.L,) * 3, self.np_dtype)
        v = np.empty_like(u)
        lib().gso_gather(self._h, u.ctypes.data, v.ctypes.data)
        return u, v


def max_threads() -> int:
    return lib().gso_max_threads()


def set_threads(n: int) -> None:
    lib().gso_set_threads(n)
