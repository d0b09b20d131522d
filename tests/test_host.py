"""CPU tests of the boundary and the host-side mirror: the C-ABI library loads and exports every
symbol include/gsb200.h declares, Settings/TOML/domain logic follow the reference, and the
multi-process plumbing works over gloo with world_size 2.  No compute call needs a GPU here."""
import ctypes as C
import os
import re
import subprocess
import sys
import textwrap

import numpy as np
import pytest

from conftest import PHYS, ROOT


def test_library_exports_every_declared_symbol(pkg):
    header = open(os.path.join(ROOT, "include", "gsb200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(gsb_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 30
    lib = C.CDLL(pkg.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in gsb200.h but not exported"
    assert declared == set(pkg.SYMBOLS), declared ^ set(pkg.SYMBOLS)
    assert pkg._lib.lib().gsb_version() == 100


def test_header_is_valid_c_and_links(pkg, tmp_path):
    """A C program that includes include/gsb200.h, takes the address of every entry point and calls the
    GPU-free ones: what a C / ccall / cgo-style binding would see."""
    exe = tmp_path / "header_check"
    libdir = os.path.dirname(pkg.LIB_PATH)
    cc = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc"
    r = subprocess.run([cc, "-std=c11", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"),
                        os.path.join(ROOT, "tests", "header_check.c"), "-o", str(exe), "-L", libdir, "-lgsb200",
                        f"-Wl,-rpath,{libdir}"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([str(exe)], capture_output=True, text=True)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    assert "38 symbols, version 100" in r.stdout
    # the Python mirror of struct gsb_config has the same size as the C one
    assert f"sizeof(gsb_config) = {C.sizeof(pkg._lib.GsbConfig)}" in r.stdout


def test_product_never_touches_the_oracle():
    """The oracle is test infrastructure: nothing under grayscott.jl_b200/ may name it."""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "grayscott.jl_b200")):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".inc")):
                text = open(os.path.join(dirpath, fn)).read()
                assert "gs_oracle" not in text.replace("oracle/gs_oracle.c holds the same definition", "")
                assert "import oracle" not in text and "from oracle" not in text


def test_settings_defaults_and_toml(sim, tmp_path):
    s = sim.Settings()
    # Structs.jl:5-27
    assert (s.L, s.steps, s.plotgap, s.F, s.k, s.dt, s.Du, s.Dv, s.noise) == (128, 20000, 200, 0.04, 0, 0.2, 0.05, 0.1, 0.0)
    assert (s.precision, s.backend, s.kernel_language) == ("Float64", "CPU", "Plain")
    toml = textwrap.dedent("""
        L = 64
        Du = 0.2
        Dv = 0.1
        F = 0.02
        k = 0.048
        dt = 1.0
        plotgap = 10
        steps = 1000
        noise = 0.1
        output = "gs-1MPI-1GPU-64L-F32.bp"
        adios_config = "adios2.yaml"
        precision = "Float32"
        backend = "B200"
    """)
    s = sim.parse_settings_toml(toml)                 # unknown key adios_config ignored (Inputs.jl:88-94)
    assert (s.L, s.steps, s.noise, s.precision, s.backend) == (64, 1000, 0.1, "Float32", "B200")
    assert not hasattr(s, "adios_config")
    assert sim.load_backend_and_lang(s) == ("b200", "plain")
    p = tmp_path / "cfg.toml"
    p.write_text(toml)
    assert sim.get_settings([str(p)]).L == 64
    with pytest.raises(ValueError, match="TOML"):     # unit-Inputs.jl:11: non-.toml -> ArgumentError
        sim.get_settings(["cfg.json"])


def test_unsupported_backend_message(sim):
    s = sim.Settings(backend="CUDA", kernel_language="KernelAbstractions")
    mcd = sim.init_domain(s, 0, 1)
    with pytest.raises(sim.GsbError, match="Backend :cuda and kernel language :kernelabstractions not supported."):
        sim.init_fields(s, mcd, "Float32")


@pytest.mark.parametrize("nranks", [1, 2, 4, 8, 6])
def test_init_domain_matches_oracle(sim, orc, nranks):
    L = 24
    s = sim.Settings(L=L)
    o = orc.Oracle(L, "Float32", dims=orc.dims_create(nranks), **PHYS)
    for r in range(nranks):
        mcd = sim.init_domain(s, r, nranks)
        info = o.rank_info(r)
        assert mcd.dims == o.dims and mcd.coords == info["coords"]
        assert mcd.proc_sizes == info["sizes"] and mcd.proc_offsets == info["offsets"]
        assert tuple(mcd.proc_neighbors[n] for n in ("west", "east", "down", "up", "south", "north")) == info["neighbors"]


def test_uneven_L_is_an_error(sim):
    with pytest.raises(sim.GsbError, match="does not divide"):
        sim.init_domain(sim.Settings(L=10), 0, 3)          # dims (3,1,1)


def test_no_gpu_fails_loudly(sim):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    s = sim.Settings(L=8, backend="B200", precision="Float32")
    with pytest.raises(sim.GsbError):
        sim.init_fields(s, sim.init_domain(s, 0, 1), "Float32")


def argparse_namespace(gpus):
    import argparse
    return argparse.Namespace(gpus=gpus, L=64, global_L=0, scaling="weak", precision="Float32", noise=0.1, arith="native")


_WORKER = r"""
import os, sys
sys.path.insert(0, {root!r})
import torch, torch.distributed as dist
import __graft_entry__ as entry
sim = entry.load_package().simulation
rank, ws = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:{port}", rank=rank, world_size=ws)
s = sim.Settings(L=16, backend="B200", precision="Float64")
mcd = sim.init_domain(s, rank, ws)
# every rank learns every rank's view; neighbour relations must be mutual and faces must match
views = [None] * ws
dist.all_gather_object(views, dict(rank=rank, sizes=mcd.proc_sizes, offsets=mcd.proc_offsets, nbr=mcd.proc_neighbors))
opp = dict(west="east", east="west", down="up", up="down", south="north", north="south")
for v in views:
    for name, q in v["nbr"].items():
        if q >= 0:
            assert views[q]["nbr"][opp[name]] == v["rank"]
assert mcd.dims == (2, 1, 1) and mcd.proc_sizes == (8, 16, 16)
# the 128-byte id travels like broadcast_unique_id's payload
t = torch.arange(128, dtype=torch.uint8) if rank == 0 else torch.zeros(128, dtype=torch.uint8)
dist.broadcast(t, src=0)
assert bytes(t.tolist()) == bytes(range(128))
# weak-sharded throughput bookkeeping of bench.py: max over ranks
ms = torch.tensor([1.0 + rank], dtype=torch.float64)
dist.all_reduce(ms, op=dist.ReduceOp.MAX)
assert ms.item() == float(ws)
dist.barrier()
dist.destroy_process_group()
print("worker", rank, "ok")
"""


def test_gloo_world_size_2(tmp_path):
    port = 29500 + os.getpid() % 2000
    script = tmp_path / "worker.py"
    script.write_text(_WORKER.format(root=ROOT, port=port))
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    for p in procs:
        out, _ = p.communicate(timeout=240)
        assert p.returncode == 0, out
        assert "ok" in out


@pytest.mark.parametrize("n", [1, 2, 3, 4, 6, 8, 12, 16, 27, 30])
def test_bench_dims_create_matches_library_and_oracle(sim, orc, n):
    """bench.py computes the decomposition itself so that neither arm needs the other's library."""
    import bench
    assert bench.dims_create(n) == sim.dims_create(n) == orc.dims_create(n)


_REF_ARM = r"""
import os, runpy, sys
sys.argv = ["bench.py", "--impl", "reference", "--steps", "2", "--warmup", "1", "--L", "64", "--gpus", "{gpus}"]
runpy.run_path({bench!r}, run_name="__main__")
maps = open("/proc/self/maps").read()
assert "libgs_oracle" in maps, "the CPU arm must have run the oracle"
assert "libgsb200" not in maps, "the CPU arm loaded the product library"
assert "grayscott_jl_b200" not in sys.modules, "the CPU arm imported the product package"
"""


@pytest.mark.parametrize("gpus,launcher_env", [(1, {}), (4, {"OMP_NUM_THREADS": "1", "RANK": "0", "WORLD_SIZE": "4"})])
def test_bench_reference_arm_prints_one_json_line(tmp_path, gpus, launcher_env):
    """bench.py --impl reference (the CPU arm the driver runs beside ours): exactly one JSON line on
    stdout with the contract's keys; the process never loads the product library; under
    torch.distributed.run's OMP_NUM_THREADS=1 it still uses every core and the same sample."""
    import json
    script = tmp_path / "ref_arm.py"
    script.write_text(_REF_ARM.format(gpus=gpus, bench=os.path.join(ROOT, "bench.py")))
    r = subprocess.run([sys.executable, str(script)], capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, **launcher_env))
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, r.stdout
    d = json.loads(lines[0])
    assert d["n_gpus"] == gpus and d["cpu_baseline"]["cores"] == len(os.sched_getaffinity(0))
    assert "L=64 cube" in d["cpu_baseline"]["sample"] and "cpu_sample_L" not in d["config"]
    import bench
    a = argparse_namespace(gpus)
    assert d["config"] == bench.workload_config(a, gpus)           # the own arm prints exactly this config
    for key in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
                "scaling", "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert key in d, key
    assert d["impl"] == "reference" and d["unit"] == "GLUPS" and d["value"] > 0 and d["vs_baseline"] is None
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert "workload" in d["config"]
