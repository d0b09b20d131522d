#!/usr/bin/env python
"""bench.py -- GLUPS (cell-updates/s) of the Gray-Scott hot path on B200, and the CPU reference arm.

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (one rank per GPU)
  python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU solver (oracle port), rank 0 only

A "step" is one iterate! (halo exchange -> fused Laplacian + reaction (+noise) -> swap) of the whole
grid.  N = 1 runs BASELINE.json's headline config: L = 1024, Float32, the physics of
examples/settings-files.toml (noise = 0.1).  N > 1 runs one 1024^3 subdomain per GPU on the
reference's Cartesian decomposition (2 -> (2,1,1), 4 -> (2,2,1), 8 -> (2,2,2) = L 2048), faces over NCCL:
weak scaling.  `--scaling strong --global-L L` splits ONE L^3 grid as the reference does
(proc_sizes = L ./ dims, communication.jl:73): BASELINE configs 4 (L = 768 Float64) and 5 (L = 2048).
Prints ONE JSON line (rank 0).

The timed region is W warm-up steps, then exactly K steps bracketed by a barrier and a device synchronise;
device time from CUDA events on the update stream, max over ranks: that is `value`.  `--windows` - 1 more
windows of K steps follow back to back and are listed beside it (per rank), to show run-to-run spread,
stragglers and what a longer run does to the clocks.

Outside the timed region the line also carries:
  parity   (N > 1)  a small noisy run over the same NCCL transport, gathered and compared bit for bit
                    with a single-GPU run of the same global grid and with the CPU oracle's N-rank run
  halo     (N > 1)  pack / transport / unpack device times per piece of the exchange, summed per step
  strong   (N > 1)  BASELINE config 4 (L = 768 Float64 split over the N GPUs), a short run
  value_reference_arith (N = 1)  the same workload in the bit-exact reference arithmetic
  cpu_baseline (N = 1), e2e (host buffers through the C ABI)
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PHYS = dict(Du=0.2, Dv=0.1, F=0.02, k=0.048, dt=1.0)      # examples/settings-files.toml:2-6
BYTES_PER_CELL = {"Float32": 16, "Float64": 32}            # read u, v + write u', v'  
NVLINK_GBS = 770.0                                         # measured peer copy, one direction (B200_PROFILING.md)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--windows", type=int, default=3, help="windows of --steps steps each: the first is the timed region, the others are listed beside it")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--L", type=int, default=1024, help="weak scaling: edge of the per-GPU subdomain")
    ap.add_argument("--global-L", type=int, default=0, help="strong scaling: edge of the global cube (split L ./ dims)")
    ap.add_argument("--precision", default="Float32", choices=["Float32", "Float64"])
    ap.add_argument("--noise", type=float, default=0.1)
    ap.add_argument("--arith", default="native", choices=["native", "reference"],
                    help="Float32 arithmetic: native FMAs (default, within 1e-5) or the reference's promotion (bit-exact)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-strong", action="store_true")
    ap.add_argument("--no-ref-arith", action="store_true")
    a = ap.parse_args()
    if a.scaling == "strong" and a.global_L <= 0:
        ap.error("--scaling strong needs --global-L")
    return a


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def mem_available_bytes():
    try:
        for ln in open("/proc/meminfo"):
            if ln.startswith("MemAvailable:"):
                return int(ln.split()[1]) * 1024
    except OSError:
        pass
    return None


def ncu_traffic(workload_key):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the update kernel, from the committed
    ncu --set full capture (profiles/traffic.json), or None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        return json.load(open(p)).get(workload_key)
    return None


def dims_create(n):
    """MPI.Dims_create(n, [0,0,0]) (communication.jl:64): prime factors, largest first, each to the
    currently smallest dimension; dims non-increasing.  Pure Python so that neither arm needs the other's
    library for it (tests/test_host.py checks it against gsb_dims_create and the oracle's)."""
    f, m, p = [], n, 2
    while p * p <= m:
        while m % p == 0:
            f.append(p)
            m //= p
        p += 1
    if m > 1:
        f.append(m)
    d = [1, 1, 1]
    for q in reversed(f):
        d[d.index(min(d))] *= q
    return tuple(sorted(d, reverse=True))


def subdomain_sizes(a, n):
    """(global extent, per-rank sizes) of the workload at n GPUs."""
    dims = dims_create(n)
    if a.scaling == "strong":
        g = (a.global_L,) * 3
        for d in dims:
            if a.global_L % d:
                raise SystemExit(f"global L = {a.global_L} does not divide by dims {dims} (communication.jl:73)")
        return g, tuple(a.global_L // d for d in dims), dims
    return tuple(a.L * d for d in dims), (a.L,) * 3, dims


def workload_config(a, n):
    g, s, dims = subdomain_sizes(a, n)
    per_gpu_gb = 4 * (s[0] + 2) * (s[1] + 2) * (s[2] + 2) * (4 if a.precision == "Float32" else 8) / 1e9
    return {
        "workload": (f"Gray-Scott iterate! (halo exchange + fused 7-pt Laplacian/reaction + noise), "
                     f"{s[0]}x{s[1]}x{s[2]} cells per GPU, global {g[0]}x{g[1]}x{g[2]}, {a.precision}, noise={a.noise}, "
                     f"Du=0.2 Dv=0.1 F=0.02 k=0.048 dt=1.0 (examples/settings-files.toml)"),
        "scaling_mode": a.scaling, "per_gpu": list(s), "global": list(g), "dims": list(dims),
        "precision": a.precision, "noise": a.noise,
        "arith": a.arith if a.precision == "Float32" else "reference",
        "l2": (f"fields are {per_gpu_gb:.1f} GB per GPU, " +
               ("far larger than the 126 MB L2; no flush needed" if per_gpu_gb > 1.0 else "L2-RESIDENT: not a roofline case")),
    }


class ClockSampler:
    """SM clock and throttle reasons DURING the timed region (NVML, 10 ms period)."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
            "hw_power_brake": getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80),
        }
        get = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = get(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.01)

    def __enter__(self):
        if self.nv:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thr:
            self._thr.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference's CPU-threaded solver, restated (oracle/gs_oracle.c; no Julia in this image).
# This function and run_reference are the only places the oracle is TIMED; it never touches the product.
# ------------------------------------------------------------------------------------------------
def host_threads():
    """Cores this process may use.  torch.distributed.run exports OMP_NUM_THREADS=1 to its workers, which
    would cripple the CPU arm: the reference runs Threads.@threads over all cores (Simulation_CPU.jl:92)."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def time_cpu_solver(L_target, precision, noise, steps, warmup, budget_s):
    """`steps` timed steps of the oracle (one xoshiro256++ draw per cell like Simulation_CPU.jl:103, all
    host threads over z-planes) on a cube of the workload: L_target itself when (warmup + steps) steps fit
    the time budget and host memory, else the largest smaller cube that does.  The choice depends on the
    host only, never on the number of GPUs.  Returns (glups, L, steps, threads, seconds)."""
    import __graft_entry__ as entry
    orc = entry.load_oracle()
    orc.build()
    threads = host_threads()
    orc.set_threads(threads)
    o = orc.Oracle(128, precision, noise=noise, noise_mode=orc.NOISE_XOSHIRO, seed=1, **PHYS)
    o.step(1)
    t0 = time.perf_counter()
    o.step(3)
    per_cell = (time.perf_counter() - t0) / 3 / 128 ** 3
    del o
    es = 4 if precision == "Float32" else 8
    avail = mem_available_bytes()
    cands = sorted({c for c in (32, 64, 128, 192, 256, 384, 512, 768, 1024) if c <= L_target} | {L_target})
    L = cands[0]
    for cand in cands:
        fits_time = per_cell * cand ** 3 * (steps + warmup) <= budget_s
        fits_mem = avail is None or 4 * (cand + 2) ** 3 * es <= 0.5 * avail
        if fits_time and fits_mem:
            L = cand
    o = orc.Oracle(L, precision, noise=noise, noise_mode=orc.NOISE_XOSHIRO, seed=1, **PHYS)
    o.step(max(warmup, 1))
    t0 = time.perf_counter()
    o.step(steps)
    dt = time.perf_counter() - t0
    return L ** 3 * steps / dt / 1e9, L, steps, threads, dt


def run_reference(a):
    """--impl reference: K timed steps of the CPU solver on a bounded sample of the same workload.  The
    host is the same one node whatever --gpus says, so the value does not depend on N."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    L_target = a.global_L if a.scaling == "strong" else a.L
    glups, L, steps, threads, dt = time_cpu_solver(L_target, a.precision, a.noise, a.steps, a.warmup, budget_s=150.0)
    sample = (f"L={L} cube of the workload, {steps} timed steps after {max(a.warmup, 1)} warm-up, {threads} OpenMP threads over "
              f"z-planes (all cores of this host; the launcher's OMP_NUM_THREADS is overridden), one xoshiro256++ draw per "
              f"cell (Simulation_CPU.jl:103); C restatement of the Julia solver (Julia itself is unavailable on this box); "
              f"the same host and sample for every --gpus N")
    line = {
        "impl": "reference", "metric": "GLUPS (cell-updates/s)", "value": glups, "unit": "GLUPS", "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": dt / steps * 1e3, "higher_is_better": True,
        "scaling": a.scaling, "vs_baseline": None, "dtype": "f32" if a.precision == "Float32" else "f64",
        "data": "synthetic (reference initial condition, Simulation_CPU.jl:14-65)",
        "config": workload_config(a, a.gpus),
        "cpu_baseline": {"value": glups, "unit": "GLUPS", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": glups, "unit": "GLUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
class Ctx:
    """Process-wide state of the GPU arm."""

    def __init__(self):
        import torch
        import __graft_entry__ as entry
        self.torch = torch
        self.pkg = entry.load_package()
        self.sim, self.lib = self.pkg.simulation, self.pkg._lib
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist = dist

    def barrier(self, *fields):
        for f in fields:
            f.sync()
        self.torch.cuda.synchronize()
        if self.dist:
            self.dist.barrier()

    def max_over_ranks(self, x):
        if not self.dist:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return t.item()

    def sum_over_ranks(self, x):
        if not self.dist:
            return x
        t = self.torch.tensor([x], dtype=self.torch.int64, device="cuda")
        self.dist.all_reduce(t)
        return int(t.item())

    def gather_list(self, x):
        if not self.dist:
            return [x]
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        out = [self.torch.zeros_like(t) for _ in range(self.world)]
        self.dist.all_gather(out, t)
        return [o.item() for o in out]


def make_fields(cx, glob, sizes, precision, noise, arith, seed=None):
    """Settings + MPICartDomain + Fields of this rank for a global box `glob` split into `sizes` per rank on
    the reference's decomposition; connects the NCCL transport when there is more than one rank."""
    sim, lib = cx.sim, cx.lib
    s = sim.Settings(L=glob[0], precision=precision, backend="B200", noise=noise, **PHYS)
    cfg = lib.GsbConfig()
    # dims, coords, neighbours of the Cartesian communicator (init_domain, communication.jl:59-96)
    sim.check(lib.lib().gsb_fill_domain(C.byref(cfg), glob[0], cx.world, cx.rank))
    mcd = sim.MPICartDomain(rank=cx.rank, nranks=cx.world, dims=tuple(cfg.dims), coords=tuple(cfg.coords),
                            proc_sizes=tuple(sizes), proc_offsets=tuple(n * c for n, c in zip(sizes, cfg.coords)),
                            proc_neighbors=dict(zip(lib.NEIGHBOR_NAMES, cfg.neighbors)))
    ar = lib.GSB_ARITH_NATIVE if (arith == "native" and precision == "Float32") else lib.GSB_ARITH_REFERENCE
    kw = {} if seed is None else {"seed": seed}
    f = sim.init_fields(s, mcd, precision, device=cx.local, arith=ar, global_extent=tuple(glob), **kw)
    if cx.world > 1:
        sim.connect_nccl(f, sim.broadcast_unique_id(cx.rank))
    return s, mcd, f


def timed_windows(cx, f, s, mcd, steps, warmup, windows, sample_clocks=False):
    """W warm-up steps, then `windows` windows of exactly `steps` steps; per window the device time of this
    rank (CUDA events on the update stream).  Returns (per-window ms of this rank, launches per window,
    clock summaries of window 0 and of the later windows)."""
    torch, sim = cx.torch, cx.sim
    stream = torch.cuda.ExternalStream(f.compute_stream(), device=cx.local)
    sim.iterate(f, s, mcd, nsteps=max(warmup, 3))
    cx.barrier(f)
    ms, launches, clocks = [], 0, [None, None]
    later = ClockSampler(cx.local) if sample_clocks else None
    for w in range(windows):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = f.launch_count()
        clk = ClockSampler(cx.local) if (sample_clocks and w == 0) else None
        if clk:
            clk.__enter__()
        elif later and w == 1:
            later.__enter__()
        e0.record(stream)
        sim.iterate(f, s, mcd, nsteps=steps)
        e1.record(stream)
        cx.barrier(f)
        if clk:
            clk.__exit__()
            clocks[0] = clk.summary()
        ms.append(e0.elapsed_time(e1))
        launches = f.launch_count() - l0
    if later and windows > 1:
        later.__exit__()
        clocks[1] = later.summary()
    return ms, launches, clocks


def halo_report(cx, f, s, mcd, nsteps=5):
    """Device times of the exchange, piece by piece, averaged over `nsteps` single steps (rank 0's view plus
    the max over ranks of the per-step sums).  The reference quantity is the 12 blocking MPI.Sendrecv! of
    communication.jl:146-177."""
    sim, lib = cx.sim, cx.lib
    f.timers_enable(True)
    n = 17
    acc = None
    phases = None
    for _ in range(nsteps):
        sim.iterate(f, s, mcd, nsteps=1)
        pk, wr, up = (C.c_double * n)(), (C.c_double * n)(), (C.c_double * n)()
        by, cnt = (C.c_int64 * n)(), C.c_int()
        sim.check(lib.lib().gsb_halo_timers(f.handle, n, pk, wr, up, by, C.byref(cnt)))
        cur = [[pk[i], wr[i], up[i], by[i]] for i in range(cnt.value)]
        acc = cur if acc is None else [[x + y for x, y in zip(p, q)] for p, q in zip(acc, cur)]
        phases = f.timers()
    f.timers_enable(False)
    pieces = [{"pack_ms": p[0] / nsteps, "wire_ms": p[1] / nsteps, "unpack_ms": p[2] / nsteps, "bytes_out": int(p[3] / nsteps),
               "wire_gbs": (p[3] / nsteps) / max(p[1] / nsteps, 1e-9) / 1e6} for p in (acc or [])]
    tot = {k: sum(p[k] for p in pieces) for k in ("pack_ms", "wire_ms", "unpack_ms")}
    bytes_out = sum(p["bytes_out"] for p in pieces)
    total_ms = sum(tot.values())
    return {
        "ms_per_step": total_ms, "pack_ms": tot["pack_ms"], "wire_ms": tot["wire_ms"], "unpack_ms": tot["unpack_ms"],
        "ms_per_step_max_over_ranks": cx.max_over_ranks(total_ms),
        "pack_unpack_ms_max_over_ranks": cx.max_over_ranks(tot["pack_ms"] + tot["unpack_ms"]),
        "bytes_out_per_gpu": bytes_out,
        "wire_gbs": bytes_out / max(tot["wire_ms"], 1e-9) / 1e6, "nvlink_peak_gbs_one_direction": NVLINK_GBS,
        "wire_frac_of_nvlink": bytes_out / max(tot["wire_ms"], 1e-9) / 1e6 / NVLINK_GBS,
        "ideal_wire_ms_at_peak": bytes_out / NVLINK_GBS / 1e6,
        "pieces": pieces, "interior_ms": phases["interior_ms"], "shell_ms": phases["shell_ms"], "step_ms": phases["step_ms"],
        "what": ("rank 0, mean of %d single steps: sums over the pieces of one step's exchange (one piece per z chunk of the "
                 "pipelined scheme); wire = the ncclSend/ncclRecv group, which includes waiting until the neighbour has "
                 "reached the same piece, so wire_gbs is a lower bound of the link rate; the exchange runs one step ahead on "
                 "its own stream, hidden under the update" % nsteps),
    }


def parity_report(cx, L=192, steps=12):
    """Outside any timed region: a noisy run of a small global grid over THIS job's NCCL transport, gathered
    to rank 0 and compared bit for bit (a) with a single-GPU run of the same global grid (north_star's halo
    criterion) and (b) with the CPU oracle's N-rank run (oracle = checker only)."""
    import numpy as np
    torch, sim = cx.torch, cx.sim
    if any(L % d for d in dims_create(cx.world)):
        return {"ok": None, "skipped": f"L = {L} does not divide by dims {dims_create(cx.world)}"}
    cases = []
    ok_all = True
    for precision, arith, halo in (("Float32", "reference", "pipeline"), ("Float32", "reference", "split"),
                                   ("Float32", "native", "pipeline"), ("Float64", "reference", "pipeline")):
        os.environ["GSB_HALO"] = halo
        s, mcd, f = make_fields(cx, (L, L, L), tuple(L // d for d in dims_create(cx.world)), precision, 0.1, arith, seed=9)
        sim.iterate(f, s, mcd, nsteps=steps)
        u, v = sim.get_fields(f)
        f.close()
        tu, tv = torch.from_numpy(u).cuda(), torch.from_numpy(v).cuda()
        gu = [torch.empty_like(tu) for _ in range(cx.world)]
        gv = [torch.empty_like(tv) for _ in range(cx.world)]
        cx.dist.all_gather(gu, tu)
        cx.dist.all_gather(gv, tv)
        case = {"precision": precision, "arith": arith, "halo": halo}
        if cx.rank == 0:
            dims = dims_create(cx.world)
            sz = tuple(L // d for d in dims)
            U = np.empty((L, L, L), u.dtype)
            V = np.empty_like(U)
            for r in range(cx.world):
                cxr, cyr, czr = r // (dims[1] * dims[2]), (r // dims[2]) % dims[1], r % dims[2]   # row-major coords
                sl = (slice(czr * sz[2], (czr + 1) * sz[2]), slice(cyr * sz[1], (cyr + 1) * sz[1]),
                      slice(cxr * sz[0], (cxr + 1) * sz[0]))
                U[sl], V[sl] = gu[r].cpu().numpy(), gv[r].cpu().numpy()
            # (a) one subdomain, one GPU, no exchange at all
            s1 = sim.Settings(L=L, precision=precision, backend="B200", noise=0.1, **PHYS)
            mcd1 = sim.init_domain(s1, 0, 1)
            ar = cx.lib.GSB_ARITH_NATIVE if arith == "native" else cx.lib.GSB_ARITH_REFERENCE
            one = sim.init_fields(s1, mcd1, precision, device=cx.local, arith=ar, seed=9)
            sim.iterate(one, s1, mcd1, nsteps=steps)
            u1, v1 = sim.get_fields(one)
            one.close()
            case["vs_single_gpu_bit_exact"] = bool(np.array_equal(U, u1) and np.array_equal(V, v1))
            # (b) the oracle's in-process N-rank run (reference arithmetic only)
            if arith == "reference":
                import __graft_entry__ as entry
                orc = entry.load_oracle()
                orc.build()
                orc.set_threads(host_threads())
                o = orc.Oracle(L, precision, dims=dims, noise=0.1, seed=9, **PHYS)
                o.step(steps)
                uo, vo = o.gather()
                case["vs_oracle_nrank_bit_exact"] = bool(np.array_equal(U, uo) and np.array_equal(V, vo))
            ok_all = ok_all and all(v for k, v in case.items() if k.startswith("vs_"))
        cases.append(case)
    os.environ.pop("GSB_HALO", None)
    return {"ok": ok_all, "L": L, "steps": steps, "noise": 0.1, "n_ranks": cx.world, "transport": "nccl", "cases": cases,
            "what": ("gather of the N-rank NCCL run == single-GPU run of the same global grid, bit for bit, and == the CPU "
                     "oracle's N-rank run (reference arithmetic); outside the timed region")}


def strong_report(cx, L=768, precision="Float64", steps=30):
    """BASELINE config 4 at this N: ONE L^3 grid split L ./ dims (communication.jl:73), a short run."""
    dims = dims_create(cx.world)
    if any(L % d for d in dims):
        return {"skipped": f"L = {L} does not divide by dims {dims}"}
    sizes = tuple(L // d for d in dims)
    s, mcd, f = make_fields(cx, (L, L, L), sizes, precision, 0.1, "reference")
    ms, launches, _ = timed_windows(cx, f, s, mcd, steps, 5, 3)
    per_win = [cx.max_over_ranks(m) for m in ms]
    halo = halo_report(cx, f, s, mcd, nsteps=3)
    kern = get_kernel(cx, f)
    f.close()
    med = statistics.median(per_win) / steps
    peak, _ = peaks()
    cells = float(L) ** 3
    return {"config": f"L={L} {precision} noise=0.1 split {dims[0]}x{dims[1]}x{dims[2]} -> {sizes[0]}x{sizes[1]}x{sizes[2]} per GPU",
            "value": cells / med / 1e6, "unit": "GLUPS", "ms_per_step": med, "steps": steps,
            "windows_ms_per_step": [m / steps for m in per_win],
            "roofline_frac_per_gpu": cells / cx.world * BYTES_PER_CELL[precision] / (med * 1e-3) / 1e9 / peak,
            "kernel": kern, "gpu_launches_per_step": cx.sum_over_ranks(launches) / steps,
            "halo": {k: halo[k] for k in ("ms_per_step", "pack_ms", "wire_ms", "unpack_ms", "bytes_out_per_gpu", "wire_gbs")}}


def get_kernel(cx, f):
    kern = [C.c_int(), C.c_int(), C.c_int(), C.c_int()]
    cx.sim.check(cx.lib.lib().gsb_get_kernel(f.handle, *[C.byref(k) for k in kern]))
    return f"stencil_{'l1' if kern[0].value == 1 else 'tma'}_kernel ty={kern[1].value} stages={kern[2].value} zchunk={kern[3].value}"


def e2e_report(cx, a, f, sizes, cells):
    """The same metric end to end through the public C ABI with HOST buffers (pinned), copies timed.

    e2e            the reference driver's own loop (GrayScott.main, src/GrayScott.jl:81-96, plotgap = 10 of
                   examples/settings-files.toml) with the calls the Julia glue makes: gsb_set_ghosted(u, v) once,
                   then per step gsb_step(1) and every plotgap steps the BLOCKING gsb_get_fields that
                   Simulation.get_fields binds (IO.write_step!, IO.jl:84-85).
    e2e_async      the same loop with gsb_get_fields_async / gsb_output_wait (B200Backend.get_fields_async!,
                   opt-in): the D2H of output k overlaps the steps up to output k+1.
    e2e_every_step upload both ghosted fields and download both results on EVERY step (no real driver does)."""
    import numpy as np
    torch, sim, lib = cx.torch, cx.sim, cx.lib
    T = np.float32 if a.precision == "Float32" else np.float64
    tdt = torch.float32 if T == np.float32 else torch.float64
    sx, sy, sz = sizes
    n_g, n_i = (sx + 2) * (sy + 2) * (sz + 2), sx * sy * sz
    es = 4 if T == np.float32 else 8
    need = 2 * n_g * es * int(os.environ.get("LOCAL_WORLD_SIZE", cx.world))
    avail = mem_available_bytes()
    skip = None
    if avail is not None and need > 0.4 * avail:
        skip = f"e2e skipped: {need / 1e9:.0f} GB of pinned host memory needed, {avail / 1e9:.0f} GB available"
    if cx.dist:                                    # all ranks take the same decision
        ok = torch.tensor([0 if skip else 1], dtype=torch.int32, device="cuda")
        cx.dist.all_reduce(ok, op=cx.dist.ReduceOp.MIN)
        if ok.item() == 0 and skip is None:
            skip = "e2e skipped: another rank lacks host memory"
    if skip:
        return {"e2e": {"value": None, "unit": "GLUPS", "h2d_bytes_per_step": None, "d2h_bytes_per_step": None, "what": skip}}
    hu = torch.empty(n_g, dtype=tdt, pin_memory=True)
    hv = torch.empty(n_g, dtype=tdt, pin_memory=True)
    L = lib.lib()
    plotgap = 10

    def refill():
        sim.check(L.gsb_get_ghosted(f.handle, lib.GSB_U, hu.data_ptr()))
        sim.check(L.gsb_get_ghosted(f.handle, lib.GSB_V, hv.data_ptr()))

    def driver_loop(k, use_async):
        sim.check(L.gsb_set_ghosted(f.handle, lib.GSB_U, hu.data_ptr()))
        sim.check(L.gsb_set_ghosted(f.handle, lib.GSB_V, hv.data_ptr()))
        n_out = 0
        for step in range(1, k + 1):
            sim.check(L.gsb_step(f.handle, 1))
            if step % plotgap == 0:
                if use_async:
                    sim.check(L.gsb_get_fields_async(f.handle, hu.data_ptr(), hv.data_ptr()))
                else:
                    sim.check(L.gsb_get_fields(f.handle, hu.data_ptr(), hv.data_ptr()))
                n_out += 1
        if use_async:
            sim.check(L.gsb_output_wait(f.handle))
        return n_out

    out = {}
    refill()
    for name, use_async in (("e2e", False), ("e2e_async", True)):
        driver_loop(plotgap, use_async)          # warm-up: staging buffers, host pages
        refill()                                 # the timed run uploads real ghosted fields again
        cx.barrier(f)
        t0 = time.perf_counter()
        n_out = driver_loop(a.steps, use_async)
        cx.barrier(f)
        dt = cx.max_over_ranks(time.perf_counter() - t0)
        h2d, d2h = 2 * n_g * es * cx.world, 2 * n_i * es * cx.world * n_out
        out[name] = {
            "value": cells * a.steps / dt / 1e9, "unit": "GLUPS", "h2d_bytes_per_step": h2d / a.steps,
            "d2h_bytes_per_step": d2h / a.steps, "steps": a.steps, "outputs": n_out, "plotgap": plotgap,
            "ms_per_step": dt / a.steps * 1e3, "pcie_gbs_per_gpu": (h2d + d2h) / cx.world / dt / 1e9,
            "what": ("GrayScott.main's loop through the C ABI with pinned host buffers: gsb_set_ghosted(u, v) once, then per "
                     "step gsb_step(1) and every plotgap = 10 steps " +
                     ("gsb_get_fields_async (+ gsb_output_wait at the end)" if use_async else
                      "the blocking gsb_get_fields, the call the Julia glue's get_fields makes for IO.write_step!") +
                     "; all inside the timed region; PCIe-bound")}
        refill()

    def one():
        sim.check(L.gsb_set_ghosted(f.handle, lib.GSB_U, hu.data_ptr()))
        sim.check(L.gsb_set_ghosted(f.handle, lib.GSB_V, hv.data_ptr()))
        sim.check(L.gsb_step(f.handle, 1))
        sim.check(L.gsb_get_fields(f.handle, hu.data_ptr(), hv.data_ptr()))

    one()
    refill()
    cx.barrier(f)
    t0 = time.perf_counter()
    for _ in range(a.e2e_steps):
        one()
    cx.barrier(f)
    dt = cx.max_over_ranks(time.perf_counter() - t0)
    out["e2e_every_step"] = {
        "value": cells * a.e2e_steps / dt / 1e9, "unit": "GLUPS", "h2d_bytes_per_step": 2 * n_g * es * cx.world,
        "d2h_bytes_per_step": 2 * n_i * es * cx.world, "steps": a.e2e_steps, "ms_per_step": dt / a.e2e_steps * 1e3,
        "pcie_gbs_per_gpu": (2 * n_g + 2 * n_i) * es * a.e2e_steps / dt / 1e9,
        "what": "per step: gsb_set_ghosted(u), gsb_set_ghosted(v) from pinned host, gsb_step(1), gsb_get_fields to pinned host"}
    return out


def run_b200(a):
    cx = Ctx()
    if cx.world != a.gpus:
        raise SystemExit(f"--gpus {a.gpus} but WORLD_SIZE={cx.world}: launch with torch.distributed.run --nproc-per-node {a.gpus}")
    glob, sizes, dims = subdomain_sizes(a, cx.world)
    s, mcd, f = make_fields(cx, glob, sizes, a.precision, a.noise, a.arith)
    ms, launches, clocks = timed_windows(cx, f, s, mcd, a.steps, a.warmup, max(a.windows, 1), sample_clocks=True)
    win_rank = [cx.gather_list(m / a.steps) for m in ms]                    # [window][rank] ms per step
    win_ms = [max(w) for w in win_rank]                                     # max over ranks, per window
    # `value` is the contract's measurement: exactly K steps right after the W warm-up steps (window 0).  The
    # further windows follow back to back and show what a longer run does (with noise on, the update kernel
    # reaches the board's power cap after ~0.1 s and the SM clock drops).
    mid = 0
    ms_step = win_ms[mid]
    launches = cx.sum_over_ranks(launches)
    cells = float(sizes[0]) * sizes[1] * sizes[2] * cx.world
    glups = cells / ms_step / 1e6

    halo = halo_report(cx, f, s, mcd) if cx.world > 1 else None
    kern = get_kernel(cx, f)
    e2e = e2e_report(cx, a, f, sizes, cells) if not a.no_e2e else {}
    f.close()

    ref_arith = None
    if cx.world == 1 and a.precision == "Float32" and a.arith == "native" and not a.no_ref_arith:
        s2, mcd2, f2 = make_fields(cx, glob, sizes, a.precision, a.noise, "reference")
        ms2, _, _ = timed_windows(cx, f2, s2, mcd2, 10, 3, 3)
        kern2 = get_kernel(cx, f2)
        f2.close()
        m2 = statistics.median(ms2) / 10
        ref_arith = {"value": cells / m2 / 1e6, "unit": "GLUPS", "ms_per_step": m2, "steps": 10, "kernel": kern2,
                     "roofline_frac": cells * BYTES_PER_CELL[a.precision] / (m2 * 1e-3) / 1e9 / peaks()[0],
                     "what": ("GSB_ARITH_REFERENCE: the reference's mixed Float32/Float64 promotion, rounding for rounding "
                              "(bit-identical to the CPU solver); the default of the Julia glue")}

    parity = parity_report(cx) if (cx.world > 1 and not a.no_parity) else None
    strong = strong_report(cx) if (cx.world > 1 and a.scaling == "weak" and not a.no_strong) else None

    if cx.rank != 0:
        if cx.dist:
            cx.dist.destroy_process_group()
        return

    peak, peak_src = peaks()
    bpc = BYTES_PER_CELL[a.precision]
    per_gpu_cells = float(sizes[0]) * sizes[1] * sizes[2]
    # at N = 1 the update kernel is the only kernel of a step: its average launch duration is ms_step
    achieved = per_gpu_cells * bpc / (ms_step * 1e-3) / 1e9
    key = f"L{sizes[0]}_{a.precision}_{a.arith if a.precision == 'Float32' else 'reference'}_noise{a.noise:g}"
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": ncu_traffic(key) if sizes[0] == sizes[1] == sizes[2] else None, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": per_gpu_cells * bpc, "kernel": kern,
                "note": ("per-GPU algorithmic bytes / per-step device time" +
                         ("" if cx.world == 1 else "; at N>1 a step is several update launches (z chunks) plus pack/unpack and NCCL"))}
    line = {
        "metric": "GLUPS (cell-updates/s)", "value": glups, "unit": "GLUPS", "n_gpus": cx.world, "steps": a.steps,
        "warmup": max(a.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": a.scaling,
        "vs_baseline": None, "dtype": "f32" if a.precision == "Float32" else "f64",
        "data": "synthetic (reference initial condition, Simulation_CPU.jl:14-65)",
        "config": workload_config(a, cx.world), "clocks": clocks[0], "gpu_launches": launches,
        "roofline": roofline,
        "windows": {"n": len(win_ms), "steps_each": a.steps, "ms_per_step": win_ms, "min": min(win_ms),
                    "median": statistics.median(win_ms), "max": max(win_ms), "value_window": mid,
                    "ms_per_step_by_rank": win_rank[mid], "ms_per_step_by_rank_all": win_rank,
                    "clocks_later_windows": clocks[1],
                    "what": ("device time per window of K steps, max over ranks, windows back to back; value = window 0 = the K "
                             "steps that follow the warm-up (the contract's timed region); by_rank = per-rank times")},
    }
    line.update(e2e)
    if halo:
        line["halo"] = halo
    if parity:
        line["parity"] = parity
    if strong:
        line["strong"] = strong
    if ref_arith:
        line["value_reference_arith"] = ref_arith
    if cx.world == 1 and not a.no_cpu_baseline:
        L_target = a.global_L if a.scaling == "strong" else a.L
        v, Ls, st, thr, secs = time_cpu_solver(L_target, a.precision, a.noise, 10, 1, budget_s=25.0)
        line["cpu_baseline"] = {
            "value": v, "unit": "GLUPS", "cores": thr, "kind": "port",
            "sample": (f"L={Ls} cube, {st} steps in {secs:.1f} s, same physics and one noise draw per cell; C restatement of "
                       f"the reference's Julia CPU-threaded solver (Julia unavailable on this box), {thr} OpenMP threads")}
    emit(line)
    if cx.dist:
        cx.dist.destroy_process_group()


_REAL_STDOUT = None


def emit(line: dict) -> None:
    """The ONE JSON line, on the process's real stdout."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    global _REAL_STDOUT
    a = parse()
    # libraries chat on stdout (NCCL prints its version there): keep stdout for the JSON line alone
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)


if __name__ == "__main__":
    main()
