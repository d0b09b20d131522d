"""The Julia glue (julia/B200Backend.jl) cannot run here (no Julia in the image).  Next-best evidence:

* CPU: `struct GsbConfig` is parsed out of the .jl file and laid out with Julia's isbits-struct rules (C
  layout: natural alignment, NTuple = inline array); field order, widths and offsets must equal the ctypes
  mirror, which tests/header_check.c ties to the C struct; every `ccall` in the glue must name a symbol the
  library exports, with the argument count of the C prototype.
* GPU: a ctypes structure is BUILT FROM the parsed Julia definition and the glue's exact call sequence
  (gsb_create by reference -> gsb_init_fields -> gsb_step(h, 1) ... -> gsb_get_fields, and the piecewise
  exchange!/calculate!/swap!) is replayed with it and checked against the oracle."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import PHYS, ROOT

JL = os.path.join(ROOT, "julia", "B200Backend.jl")
SCALARS = {"Int32": C.c_int32, "Int64": C.c_int64, "UInt64": C.c_uint64, "Float64": C.c_double, "UInt32": C.c_uint32,
           "Cint": C.c_int}


def julia_struct_fields():
    text = open(JL).read()
    body = re.search(r"^struct GsbConfig\n(.*?)^end", text, flags=re.S | re.M).group(1)
    fields = []
    for ln in body.strip().splitlines():
        name, ty = [x.strip() for x in ln.split("::")]
        m = re.fullmatch(r"NTuple\{(\d+),\s*(\w+)\}", ty)
        fields.append((name, SCALARS[m.group(2)] * int(m.group(1)) if m else SCALARS[ty]))
    return fields


def julia_config_type():
    return type("JlGsbConfig", (C.Structure,), {"_fields_": julia_struct_fields()})


def test_julia_struct_matches_the_c_struct(pkg):
    jl = julia_config_type()
    py = pkg._lib.GsbConfig
    assert C.sizeof(jl) == C.sizeof(py)
    assert [n for n, _ in jl._fields_] == [n for n, _ in py._fields_]
    for (n, _), (m, _) in zip(jl._fields_, py._fields_):
        a, b = getattr(jl, n), getattr(py, m)
        assert (a.offset, a.size) == (b.offset, b.size), n
    # and the C header itself, field by field
    header = open(os.path.join(ROOT, "include", "gsb200.h")).read()
    body = re.search(r"typedef struct gsb_config \{(.*?)\} gsb_config;", header, flags=re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    cnames = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        ty, rest = decl.split(None, 1)
        for item in rest.split(","):
            cnames.append((ty, item.strip()))
    ctype = {"int32_t": C.c_int32, "int64_t": C.c_int64, "uint64_t": C.c_uint64, "double": C.c_double}
    assert len(cnames) == len(jl._fields_)
    for (ty, item), (n, t) in zip(cnames, jl._fields_):
        m = re.fullmatch(r"(\w+)\[(\d+)\]", item)
        cname, want = (m.group(1), ctype[ty] * int(m.group(2))) if m else (item, ctype[ty])
        assert cname.rstrip("_") == n.rstrip("_"), (cname, n)
        assert C.sizeof(want) == C.sizeof(t), n


def test_every_ccall_names_an_exported_symbol_with_the_right_arity(pkg):
    text = open(JL).read()
    calls = re.findall(r"ccall\(\(:(\w+), libgsb\), (\w+), \((.*?)\)[,)]", text)
    assert len(calls) >= 12
    protos = {name: args for name, _, args in pkg._lib._PROTOS}
    seen = set()
    for name, ret, args in calls:
        assert name in protos, f"{name} is not exported by libgsb200.so"
        nargs = len([a for a in args.split(",") if a.strip()])
        assert nargs == len(protos[name]), (name, nargs, len(protos[name]))
        assert ret in ("Cint", "Cstring")
        seen.add(name)
    for must in ("gsb_create", "gsb_init_fields", "gsb_step", "gsb_exchange", "gsb_calculate", "gsb_swap", "gsb_get_fields",
                 "gsb_get_fields_async", "gsb_output_wait", "gsb_get_ghosted", "gsb_comm_init", "gsb_get_unique_id",
                 "gsb_destroy", "gsb_last_error", "gsb_device_count"):
        assert must in seen, must


@pytest.mark.gpu
@pytest.mark.parametrize("prec", ["Float32", "Float64"])
def test_glue_call_sequence_replayed_with_the_julia_layout(pkg, orc, prec):
    """init_fields / iterate! / exchange!+calculate!+swap! / get_fields as the glue issues them, with a config
    structure laid out from the Julia definition and passed by reference (Ref{GsbConfig})."""
    L = pkg._lib.lib()
    T = np.float32 if prec == "Float32" else np.float64
    n = 24
    cfg = julia_config_type()()
    cfg.dtype, cfg.arith, cfg.L = (0 if prec == "Float32" else 1), 0, n
    for i in range(3):
        cfg.global_[i], cfg.dims[i], cfg.coords[i], cfg.sizes[i], cfg.offsets[i] = 0, 1, 0, n, 0
    for i in range(6):
        cfg.neighbors[i] = -1                                     # MPI.PROC_NULL
    cfg.Du, cfg.Dv, cfg.F, cfg.k, cfg.dt, cfg.noise = PHYS["Du"], PHYS["Dv"], PHYS["F"], PHYS["k"], PHYS["dt"], 0.1
    cfg.seed, cfg.device, cfg.rank, cfg.nranks = 0x5EED, 0, 0, 1
    h = C.c_void_p()
    create = L.gsb_create
    saved = create.argtypes
    create.argtypes = [C.POINTER(C.c_void_p), C.c_void_p]          # the glue passes Ref{GsbConfig}: a plain pointer
    try:
        assert create(C.byref(h), C.cast(C.byref(cfg), C.c_void_p)) == 0, L.gsb_last_error()
    finally:
        create.argtypes = saved
    assert L.gsb_init_fields(h) == 0
    for _ in range(3):
        assert L.gsb_step(h, 1) == 0                              # iterate!
    for _ in range(2):                                            # exchange!, calculate!, swap!
        assert L.gsb_exchange(h) == 0
        assert L.gsb_calculate(h) == 0
        assert L.gsb_swap(h) == 0
    u, v = np.empty((n, n, n), T), np.empty((n, n, n), T)
    assert L.gsb_get_fields(h, u.ctypes.data, v.ctypes.data) == 0  # get_fields for IO.write_step!
    g = np.empty((n + 2,) * 3, T)
    assert L.gsb_get_ghosted(h, 0, g.ctypes.data) == 0             # Array(fields.u)
    assert L.gsb_destroy(h) == 0
    o = orc.Oracle(n, prec, noise=0.1, noise_mode=orc.NOISE_PHILOX, seed=0x5EED, **PHYS)
    o.step(5)
    uo, vo = o.get_fields()
    assert np.array_equal(u, uo) and np.array_equal(v, vo)
    assert np.array_equal(g[1:-1, 1:-1, 1:-1], uo)
