"""Host-side mirror of GrayScott.jl's Simulation API for `backend = "B200"`.

The reference's host language is Julia, which does not exist in this image; the Julia glue a
maintainer would load is julia/B200Backend.jl.  This module is the same
interface in Python above the same C ABI (include/gsb200.h), with the reference's names,
argument meaning and error behaviour, so the parity tests read like the reference's tests:

    Settings, SettingsKeys, MPICartDomain, Fields      src/simulation/Structs.jl:4-93
    get_settings, parse_settings_toml,
    load_backend_and_lang                              src/simulation/Inputs.jl:20-120
    initialization, init_domain, exchange, finalize    src/simulation/communication.jl:15-199
    init_fields, iterate, get_fields                   src/simulation/public.jl:18-78,
                                                       src/simulation/Simulation_CPU.jl:125-133

`iterate!`/`exchange!`/`calculate!` are spelled iterate/exchange/calculate.  Only
backend "B200" with kernel_language "Plain" is implemented here; anything else raises the
reference's own message (public.jl:31-32,77-78).  Host arrays are numpy, indexed [z, y, x]
(= Julia's column-major [x, y, z] memory order).
"""
from __future__ import annotations

import ctypes as C
import dataclasses
import os
import tomllib
from typing import Optional

import numpy as np

from . import _lib
from ._lib import GsbConfig, GsbError, check, lib


# ------------------------------------------------------------------ Structs.jl
@dataclasses.dataclass
class Settings:
    """src/simulation/Structs.jl:4-28 (same keys, same defaults; `backend` default kept "CPU")."""
    L: int = 128
    steps: int = 20000
    plotgap: int = 200
    F: float = 0.04
    k: float = 0.0
    dt: float = 0.2
    Du: float = 0.05
    Dv: float = 0.1
    noise: float = 0.0
    output: str = "foo.bp"
    checkpoint: bool = False
    checkpoint_freq: int = 2000
    checkpoint_output: str = "ckpt.bp"
    restart: bool = False
    restart_input: str = "ckpt.bp"
    mesh_type: str = "image"
    precision: str = "Float64"
    backend: str = "CPU"
    kernel_language: str = "Plain"
    verbose: bool = False


SettingsKeys = frozenset(f.name for f in dataclasses.fields(Settings))  # Structs.jl:31-52

# extra, B200-only knobs carried beside Settings (not TOML keys of the reference)
DEFAULT_SEED = 0x5EED


@dataclasses.dataclass
class MPICartDomain:
    """src/simulation/Structs.jl:57-73.  cart_comm is replaced by (rank, nranks)."""
    rank: int = 0
    nranks: int = 1
    dims: tuple = (0, 0, 0)
    coords: tuple = (0, 0, 0)
    proc_sizes: tuple = (128, 128, 128)
    proc_offsets: tuple = (1, 1, 1)
    proc_neighbors: dict = dataclasses.field(
        default_factory=lambda: {n: -1 for n in _lib.NEIGHBOR_NAMES})


class B200Array:
    """Opaque device array of a Fields slot: size (sx+2, sy+2, sz+2); np.asarray() = ghosted D2H
    (what `Array(fields.u)` is for a CuArray)."""

    def __init__(self, fields: "Fields", slot: int):
        self._fields, self.slot = fields, slot

    @property
    def shape(self):
        sx, sy, sz = self._fields.sizes
        return (sx + 2, sy + 2, sz + 2)

    def __array__(self, dtype=None, copy=None):
        a = self._fields.get_ghosted(self.slot)
        return a if dtype is None else a.astype(dtype)

    def set(self, host: np.ndarray) -> None:
        self._fields.set_ghosted(self.slot, host)


class Fields:
    """src/simulation/Structs.jl:82-93: u, v, u_temp, v_temp (+ the face datatypes, which live
    inside the library as pack/unpack index maps)."""

    def __init__(self, handle, cfg: GsbConfig, np_dtype):
        self._h = handle
        self.cfg = cfg
        self.dtype = np_dtype
        self.sizes = tuple(cfg.sizes)
        self.u, self.v = B200Array(self, _lib.GSB_U), B200Array(self, _lib.GSB_V)
        self.u_temp, self.v_temp = B200Array(self, _lib.GSB_U_TEMP), B200Array(self, _lib.GSB_V_TEMP)

    # --- lifetime
    def close(self):
        if self._h:
            lib().gsb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def handle(self):
        return self._h

    # --- data access
    def get_ghosted(self, slot: int) -> np.ndarray:
        sx, sy, sz = self.sizes
        a = np.empty((sz + 2, sy + 2, sx + 2), self.dtype)
        check(lib().gsb_get_ghosted(self._h, slot, a.ctypes.data))
        return a

    def set_ghosted(self, slot: int, host: np.ndarray) -> None:
        sx, sy, sz = self.sizes
        a = np.ascontiguousarray(host, self.dtype)
        if a.shape != (sz + 2, sy + 2, sx + 2):
            raise ValueError(f"expected shape {(sz + 2, sy + 2, sx + 2)}, got {a.shape}")
        check(lib().gsb_set_ghosted(self._h, slot, a.ctypes.data))

    def sync(self):
        check(lib().gsb_sync(self._h))

    @property
    def step(self) -> int:
        s = C.c_int64()
        check(lib().gsb_get_step(self._h, C.byref(s)))
        return s.value

    @step.setter
    def step(self, value: int):
        check(lib().gsb_set_step(self._h, value))

    def stats(self, slot: int = _lib.GSB_U, thresh: float = 0.1) -> dict:
        out = (C.c_double * 5)()
        check(lib().gsb_field_stats(self._h, slot, thresh, out))
        n = float(np.prod(self.sizes))
        mean = out[0] / n
        return dict(sum=out[0], sumsq=out[1], min=out[2], max=out[3], count_above=out[4], mean=mean,
                    var=out[1] / n - mean * mean, n=n)

    def pdf(self, slot: int, nbins: int, vmin: float, vmax: float):
        """_compute_pdf (src/analysis/pdfcalc.jl:13-49): (pdf[sz, nbins] counts, bins[nbins])."""
        pdf = np.empty((self.sizes[2], nbins), np.uint32)
        bins = np.empty(nbins, np.float64)
        check(lib().gsb_pdf(self._h, slot, nbins, vmin, vmax, pdf.ctypes.data, bins.ctypes.data))
        return pdf, bins

    def set_kernel(self, kind=-1, ty=0, stages=0, zchunk=0):
        check(lib().gsb_set_kernel(self._h, kind, ty, stages, zchunk))

    def get_kernel(self):
        k = [C.c_int() for _ in range(4)]
        check(lib().gsb_get_kernel(self._h, *[C.byref(x) for x in k]))
        return tuple(x.value for x in k)

    def timers_enable(self, on=True):
        check(lib().gsb_timers_enable(self._h, int(on)))

    def timers(self) -> dict:
        out = (C.c_double * 4)()
        check(lib().gsb_timers(self._h, out))
        return dict(halo_ms=out[0], interior_ms=out[1], shell_ms=out[2], step_ms=out[3])

    def launch_count(self) -> int:
        n = C.c_int64()
        check(lib().gsb_launch_count(self._h, C.byref(n)))
        return n.value

    def compute_stream(self) -> int:
        p = C.c_void_p()
        check(lib().gsb_compute_stream(self._h, C.byref(p)))
        return p.value or 0

    def halo_bytes(self) -> int:
        b = C.c_int64()
        check(lib().gsb_halo_bytes(self._h, C.byref(b)))
        return b.value


# ------------------------------------------------------------------ Inputs.jl
def parse_settings_toml(toml_contents: str) -> Settings:
    """Inputs.jl:80-97: unknown keys are ignored silently."""
    d = tomllib.loads(toml_contents)
    s = Settings()
    for key, value in d.items():
        if key in SettingsKeys:
            setattr(s, key, value)
    return s


def get_settings(args) -> Settings:
    """Inputs.jl:20-35: exactly one positional argument, which must end in .toml."""
    if len(args) != 1:
        raise ValueError("usage: <config_file.toml>")
    config_file = args[0]
    if not config_file.endswith(".toml"):
        ext = config_file.split(".")[-1]
        raise ValueError(f"Config file must be in TOML format. Extension not recognized: {ext}\n")
    with open(config_file, "r") as f:
        return parse_settings_toml(f.read())


def load_backend_and_lang(settings: Settings):
    """Inputs.jl:110-120."""
    return settings.backend.lower(), settings.kernel_language.lower()


def _require_b200(settings: Settings):
    backend, lang = load_backend_and_lang(settings)
    if (backend, lang) != ("b200", "plain"):
        # the reference's fallback methods, public.jl:31-32,77-78
        raise GsbError(f"Backend :{backend} and kernel language :{lang} not supported.")


def _precision(T):
    """communication.jl:27: settings.precision names a Julia type."""
    if isinstance(T, str):
        try:
            return {"Float32": (np.float32, _lib.GSB_F32), "Float64": (np.float64, _lib.GSB_F64)}[T]
        except KeyError:
            raise GsbError(f"precision {T!r} is not supported by the B200 backend (Float32, Float64)")
    T = np.dtype(T)
    if T == np.float32:
        return np.float32, _lib.GSB_F32
    if T == np.float64:
        return np.float64, _lib.GSB_F64
    raise GsbError(f"element type {T} is not supported by the B200 backend")


# ------------------------------------------------------------------ communication.jl
def dims_create(nranks: int):
    d = (C.c_int32 * 3)()
    check(lib().gsb_dims_create(nranks, d))
    return tuple(d)


def init_domain(settings: Settings, rank: int = 0, nranks: int = 1) -> MPICartDomain:
    """communication.jl:59-96 without MPI: Dims_create, row-major coords, L ./ dims sizes,
    0-based offsets, non-periodic Cart_shift neighbours (-1 = PROC_NULL)."""
    cfg = GsbConfig()
    check(lib().gsb_fill_domain(C.byref(cfg), settings.L, nranks, rank))
    return MPICartDomain(rank=rank, nranks=nranks, dims=tuple(cfg.dims), coords=tuple(cfg.coords),
                         proc_sizes=tuple(cfg.sizes), proc_offsets=tuple(cfg.offsets),
                         proc_neighbors=dict(zip(_lib.NEIGHBOR_NAMES, cfg.neighbors)))


def make_config(settings: Settings, mcd: MPICartDomain, T, *, device: int = 0, seed: int = DEFAULT_SEED,
                arith: Optional[int] = None, global_extent=None) -> GsbConfig:
    _, code = _precision(T)
    cfg = GsbConfig()
    cfg.dtype = code
    cfg.arith = _lib.GSB_ARITH_REFERENCE if arith is None else arith
    cfg.L = settings.L
    g = global_extent or (settings.L,) * 3
    for i in range(3):
        cfg.global_[i] = g[i]
        cfg.dims[i] = mcd.dims[i]
        cfg.coords[i] = mcd.coords[i]
        cfg.sizes[i] = mcd.proc_sizes[i]
        cfg.offsets[i] = mcd.proc_offsets[i]
    for i, n in enumerate(_lib.NEIGHBOR_NAMES):
        cfg.neighbors[i] = mcd.proc_neighbors[n]
    cfg.Du, cfg.Dv, cfg.F, cfg.k, cfg.dt, cfg.noise = (settings.Du, settings.Dv, settings.F, settings.k,
                                                       settings.dt, settings.noise)
    cfg.seed = seed
    cfg.device, cfg.rank, cfg.nranks = device, mcd.rank, mcd.nranks
    return cfg


# ------------------------------------------------------------------ public.jl
def init_fields(settings: Settings, mcd: MPICartDomain, T, **kw) -> Fields:
    """public.jl:18-25 -> init_fields(::Val{:b200}, ::Val{:plain}, settings, mcd, T)."""
    _require_b200(settings)
    np_dtype, _ = _precision(T)
    cfg = make_config(settings, mcd, T, **kw)
    h = C.c_void_p()
    check(lib().gsb_create(C.byref(h), C.byref(cfg)))
    fields = Fields(h, cfg, np_dtype)
    check(lib().gsb_init_fields(h))
    return fields


def exchange(fields: Fields, mcd: MPICartDomain = None) -> None:
    """exchange!(fields, mcd), communication.jl:138-199."""
    check(lib().gsb_exchange(fields.handle))


def calculate(fields: Fields, settings: Settings = None, mcd: MPICartDomain = None) -> None:
    """calculate!(::Val{:b200}, ::Val{:plain}, fields, settings, mcd), Simulation_CPU.jl:77-113."""
    check(lib().gsb_calculate(fields.handle))


def swap(fields: Fields) -> None:
    """public.jl:67-68; the B200Array wrappers keep naming the same slots."""
    check(lib().gsb_swap(fields.handle))


def iterate(fields: Fields, settings: Settings, mcd: MPICartDomain = None, nsteps: int = 1) -> None:
    """iterate!(fields, settings, mcd), public.jl:45-71.  nsteps > 1 enqueues several steps at once."""
    _require_b200(settings)
    check(lib().gsb_step(fields.handle, nsteps))


def iterate_group(fields_list, settings: Settings, nsteps: int = 1) -> None:
    """All in-process ranks one step forward (the harness's stand-in for `mpirun -n N`)."""
    _require_b200(settings)
    hs = (C.c_void_p * len(fields_list))(*[f.handle for f in fields_list])
    check(lib().gsb_step_group(hs, len(fields_list), nsteps))


def connect_local(fields_list) -> None:
    hs = (C.c_void_p * len(fields_list))(*[f.handle for f in fields_list])
    check(lib().gsb_comm_init_local(hs, len(fields_list)))


def get_fields(fields: Fields, out=None):
    """get_fields(::Val{:b200}, fields), Simulation_CPU.jl:125-133: ghost-free host copies."""
    sx, sy, sz = fields.sizes
    if out is None:
        u = np.empty((sz, sy, sx), fields.dtype)
        v = np.empty_like(u)
    else:
        u, v = out
    check(lib().gsb_get_fields(fields.handle, u.ctypes.data, v.ctypes.data))
    return u, v


def get_fields_async(fields: Fields, u_host_ptr: int, v_host_ptr: int) -> None:
    """Non-blocking get_fields into caller-owned (pinned) host memory; pair with output_wait."""
    check(lib().gsb_get_fields_async(fields.handle, u_host_ptr, v_host_ptr))


def output_wait(fields: Fields) -> None:
    check(lib().gsb_output_wait(fields.handle))


def calculate_arrays(u_dev_ptr, v_dev_ptr, u_out_ptr, v_out_ptr, sizes, T, settings: Settings, *,
                     offsets=(0, 0, 0), seed=DEFAULT_SEED, step=0, arith=_lib.GSB_ARITH_REFERENCE, stream=None):
    """One calculate! on device arrays in the reference layout (raw device pointers as ints)."""
    _, code = _precision(T)
    s = (C.c_int64 * 3)(*sizes)
    o = (C.c_int64 * 3)(*offsets)
    check(lib().gsb_calculate_arrays(code, arith, s, o, u_dev_ptr, v_dev_ptr, u_out_ptr, v_out_ptr, settings.Du,
                                     settings.Dv, settings.F, settings.k, settings.dt, settings.noise, seed, step,
                                     stream))


# ------------------------------------------------------------------ initialization / finalize
def broadcast_unique_id(rank: int) -> bytes:
    """MPI.Bcast of the 128-byte NCCL id, over torch.distributed (any backend)."""
    import torch
    import torch.distributed as dist
    buf = (C.c_ubyte * 128)()
    if rank == 0:
        check(lib().gsb_get_unique_id(buf))
    t = torch.tensor(list(bytes(buf)), dtype=torch.uint8)
    if dist.get_backend() == "nccl":
        t = t.cuda()
    dist.broadcast(t, src=0)
    return bytes(t.cpu().tolist())


def connect_nccl(fields: Fields, id128: bytes) -> None:
    buf = (C.c_ubyte * 128).from_buffer_copy(id128)
    check(lib().gsb_comm_init(fields.handle, buf))


def initialization(args, **kw):
    """communication.jl:15-33: settings, (torch.distributed instead of MPI) rank/size, domain, fields.
    Returns (comm_info, settings, mcd, fields) like the reference returns (comm, settings, mcd, fields)."""
    settings = get_settings(list(args))
    rank = int(os.environ.get("RANK", "0"))
    nranks = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    mcd = init_domain(settings, rank, nranks)
    fields = init_fields(settings, mcd, settings.precision, device=local_rank, **kw)
    if nranks > 1:
        import torch.distributed as dist
        if not dist.is_initialized():
            dist.init_process_group("gloo")
        connect_nccl(fields, broadcast_unique_id(rank))
    return dict(rank=rank, nranks=nranks), settings, mcd, fields


def finalize():
    """communication.jl:40-46."""
    try:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()
    except ImportError:
        pass
