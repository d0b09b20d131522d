"""ctypes binding of libgsb200.so (include/gsb200.h).

There is no fallback: if the CUDA library has not been built the import fails loudly, and every
compute entry point fails when no B200 is present.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgsb200.so")

GSB_F32, GSB_F64 = 0, 1
GSB_ARITH_REFERENCE, GSB_ARITH_NATIVE = 0, 1
GSB_KERNEL_TMA, GSB_KERNEL_L1 = 0, 1
GSB_U, GSB_V, GSB_U_TEMP, GSB_V_TEMP = 0, 1, 2, 3
NEIGHBOR_NAMES = ("west", "east", "down", "up", "south", "north")  # communication.jl:90-92


class GsbConfig(C.Structure):
    _fields_ = [
        ("dtype", C.c_int32), ("arith", C.c_int32),
        ("L", C.c_int64), ("global_", C.c_int64 * 3),
        ("dims", C.c_int32 * 3), ("coords", C.c_int32 * 3),
        ("sizes", C.c_int64 * 3), ("offsets", C.c_int64 * 3),
        ("neighbors", C.c_int32 * 6),
        ("Du", C.c_double), ("Dv", C.c_double), ("F", C.c_double), ("k", C.c_double),
        ("dt", C.c_double), ("noise", C.c_double),
        ("seed", C.c_uint64),
        ("device", C.c_int32), ("rank", C.c_int32), ("nranks", C.c_int32),
    ]


class GsbError(RuntimeError):
    pass


# every symbol include/gsb200.h declares: (name, restype, argtypes)
_vp, _i, _i64 = C.c_void_p, C.c_int, C.c_int64
_PROTOS = [
    ("gsb_version", _i, []),
    ("gsb_last_error", C.c_char_p, []),
    ("gsb_create", _i, [C.POINTER(_vp), C.POINTER(GsbConfig)]),
    ("gsb_destroy", _i, [_vp]),
    ("gsb_dims_create", _i, [_i, C.POINTER(C.c_int32)]),
    ("gsb_fill_domain", _i, [C.POINTER(GsbConfig), _i64, _i, _i]),
    ("gsb_init_fields", _i, [_vp]),
    ("gsb_exchange", _i, [_vp]),
    ("gsb_calculate", _i, [_vp]),
    ("gsb_swap", _i, [_vp]),
    ("gsb_step", _i, [_vp, _i]),
    ("gsb_step_group", _i, [C.POINTER(_vp), _i, _i]),
    ("gsb_get_fields", _i, [_vp, _vp, _vp]),
    ("gsb_get_fields_async", _i, [_vp, _vp, _vp]),
    ("gsb_output_wait", _i, [_vp]),
    ("gsb_get_ghosted", _i, [_vp, _i, _vp]),
    ("gsb_set_ghosted", _i, [_vp, _i, _vp]),
    ("gsb_alloc_host", _i, [C.POINTER(_vp), C.c_size_t]),
    ("gsb_free_host", _i, [_vp]),
    ("gsb_sync", _i, [_vp]),
    ("gsb_get_step", _i, [_vp, C.POINTER(_i64)]),
    ("gsb_set_step", _i, [_vp, _i64]),
    ("gsb_field_stats", _i, [_vp, _i, C.c_double, C.POINTER(C.c_double)]),
    ("gsb_pdf", _i, [_vp, _i, _i, C.c_double, C.c_double, _vp, _vp]),
    ("gsb_get_unique_id", _i, [_vp]),
    ("gsb_comm_init", _i, [_vp, _vp]),
    ("gsb_comm_init_local", _i, [C.POINTER(_vp), _i]),
    ("gsb_set_kernel", _i, [_vp, _i, _i, _i, _i]),
    ("gsb_get_kernel", _i, [_vp] + [C.POINTER(_i)] * 4),
    ("gsb_timers_enable", _i, [_vp, _i]),
    ("gsb_timers", _i, [_vp, C.POINTER(C.c_double)]),
    ("gsb_halo_bytes", _i, [_vp, C.POINTER(_i64)]),
    ("gsb_halo_timers", _i, [_vp, _i, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double),
                             C.POINTER(_i64), C.POINTER(_i)]),
    ("gsb_device_count", _i, [C.POINTER(_i)]),
    ("gsb_launch_count", _i, [_vp, C.POINTER(_i64)]),
    ("gsb_compute_stream", _i, [_vp, C.POINTER(_vp)]),
    ("gsb_calculate_arrays", _i, [_i, _i, C.POINTER(_i64), C.POINTER(_i64), _vp, _vp, _vp, _vp] +
     [C.c_double] * 6 + [C.c_uint64, C.c_uint32, _vp]),
    ("gsb_field_stats_noalloc", _i, [_i, C.POINTER(_i64), _vp, C.c_double, _vp, _vp, C.c_char_p, C.c_size_t]),
]
SYMBOLS = [p[0] for p in _PROTOS]

_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise GsbError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(or make -C grayscott.jl_b200/csrc).  This backend has no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        for name, res, args in _PROTOS:
            fn = getattr(L, name)  # AttributeError if the library does not export it
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise GsbError(lib().gsb_last_error().decode(errors="replace"))
