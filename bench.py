#!/usr/bin/env python
"""bench.py -- GLUPS (cell-updates/s) of the Gray-Scott hot path on B200, and the CPU reference arm.

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (one rank per GPU)
  python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU solver (oracle port), rank 0 only

A "step" is one iterate! (halo exchange -> fused Laplacian + reaction (+noise) -> swap) of the whole
grid.  N = 1 runs BASELINE.json's headline config: L = 1024, Float32, the physics of
examples/settings-files.toml (noise = 0.1).  N > 1 runs one 1024^3 subdomain per GPU on the
reference's Cartesian decomposition (2 -> (2,1,1), 4 -> (2,2,1), 8 -> (2,2,2) = L 2048), faces over NCCL:
weak scaling.  Prints ONE JSON line (rank 0).
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


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--L", type=int, default=1024, help="edge of the per-GPU subdomain")
    ap.add_argument("--precision", default="Float32", choices=["Float32", "Float64"])
    ap.add_argument("--noise", type=float, default=0.1)
    ap.add_argument("--arith", default="native", choices=["native", "reference"],
                    help="Float32 arithmetic: native FMAs (default, within 1e-5) or the reference's promotion (bit-exact)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


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


class ClockSampler:
    """SM clock and throttle reasons DURING the timed region (NVML, 50 ms period)."""

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
            time.sleep(0.05)

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
# CPU arm: the reference's CPU-threaded solver, restated (oracle/gs_oracle.c; no Julia in this image)
# ------------------------------------------------------------------------------------------------
def cpu_solver_glups(L_target, precision, noise, budget_s, min_steps=2, max_steps=20):
    """Times the oracle (one xoshiro256++ draw per cell like Simulation_CPU.jl:103, all host threads) on
    the largest cube <= L_target whose run fits the time budget.  Returns (glups, L, steps, threads, seconds)."""
    import __graft_entry__ as entry
    orc = entry.load_oracle()
    orc.build()
    threads = orc.max_threads()
    o = orc.Oracle(128, precision, noise=noise, noise_mode=orc.NOISE_XOSHIRO, seed=1, **PHYS)
    o.step(1)
    t0 = time.perf_counter()
    o.step(3)
    per_cell = (time.perf_counter() - t0) / 3 / 128 ** 3
    del o
    L = 128
    for cand in (256, 384, 512, 768, 1024):
        if cand <= L_target and per_cell * cand ** 3 * (min_steps + 1) <= budget_s:
            L = cand
    steps = int(max(min_steps, min(max_steps, budget_s / (per_cell * L ** 3) - 1)))
    o = orc.Oracle(L, precision, noise=noise, noise_mode=orc.NOISE_XOSHIRO, seed=1, **PHYS)
    o.step(1)
    t0 = time.perf_counter()
    o.step(steps)
    dt = time.perf_counter() - t0
    return L ** 3 * steps / dt / 1e9, L, steps, threads, dt


def run_reference(a):
    """--impl reference: K timed steps of the CPU solver on a bounded sample of the same workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import __graft_entry__ as entry
    orc = entry.load_oracle()
    orc.build()
    threads = orc.max_threads()
    # size the sample so (W + K) steps end within ~2 minutes
    o = orc.Oracle(128, a.precision, noise=a.noise, noise_mode=orc.NOISE_XOSHIRO, seed=1, **PHYS)
    o.step(1)
    t0 = time.perf_counter()
    o.step(3)
    per_cell = (time.perf_counter() - t0) / 3 / 128 ** 3
    del o
    L = 64
    for cand in (128, 192, 256, 384, 512, 768, 1024):
        if cand <= a.L and per_cell * cand ** 3 * (a.steps + a.warmup) <= 120.0:
            L = cand
    o = orc.Oracle(L, a.precision, noise=a.noise, noise_mode=orc.NOISE_XOSHIRO, seed=1, **PHYS)
    o.step(a.warmup)
    t0 = time.perf_counter()
    o.step(a.steps)
    dt = time.perf_counter() - t0
    glups = L ** 3 * a.steps / dt / 1e9
    sample = (f"L={L} cube of the L={a.L} workload, {a.steps} steps, {threads} OpenMP threads over z-planes, "
              f"one xoshiro256++ draw per cell (Simulation_CPU.jl:103); C restatement of the Julia solver "
              f"(Julia itself is unavailable on this box)")
    line = {
        "impl": "reference", "metric": "GLUPS (cell-updates/s)", "value": glups, "unit": "GLUPS", "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": dt / a.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32" if a.precision == "Float32" else "f64",
        "data": "synthetic (reference initial condition)",
        "config": workload_config(a, a.gpus, cpu_sample_L=L),
        "cpu_baseline": {"value": glups, "unit": "GLUPS", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": glups, "unit": "GLUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def workload_config(a, n, cpu_sample_L=None):
    import __graft_entry__ as entry
    dims = entry.load_package().simulation.dims_create(n)
    g = [a.L * d for d in dims]
    c = {
        "workload": (f"Gray-Scott iterate! (halo exchange + fused 7-pt Laplacian/reaction + noise), "
                     f"{a.L}^3 cells per GPU, global {g[0]}x{g[1]}x{g[2]}, {a.precision}, noise={a.noise}, "
                     f"Du=0.2 Dv=0.1 F=0.02 k=0.048 dt=1.0 (examples/settings-files.toml)"),
        "L_per_gpu": a.L, "global": g, "dims": list(dims), "precision": a.precision, "noise": a.noise,
        "arith": a.arith if a.precision == "Float32" else "reference",
        "l2": "fields are 4 x 4.3 GB per GPU, far larger than the 126 MB L2; no flush needed",
    }
    if cpu_sample_L:
        c["cpu_sample_L"] = cpu_sample_L
    return c


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def run_b200(a):
    import numpy as np
    import torch
    import __graft_entry__ as entry
    pkg = entry.load_package()
    sim, lib = pkg.simulation, pkg._lib
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != a.gpus:
        raise SystemExit(f"--gpus {a.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run --nproc-per-node {a.gpus}")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    # weak scaling on the reference's decomposition: every rank owns an L^3 subdomain
    dims = sim.dims_create(world)
    g = tuple(a.L * d for d in dims)
    s = sim.Settings(L=g[0], precision=a.precision, backend="B200", noise=a.noise, **PHYS)
    cfg = lib.GsbConfig()
    sim.check(lib.lib().gsb_fill_domain(C.byref(cfg), g[0], world, rank))   # dims, coords, neighbours
    mcd = sim.MPICartDomain(rank=rank, nranks=world, dims=tuple(cfg.dims), coords=tuple(cfg.coords),
                            proc_sizes=(a.L,) * 3, proc_offsets=tuple(a.L * c for c in cfg.coords),
                            proc_neighbors=dict(zip(lib.NEIGHBOR_NAMES, cfg.neighbors)))
    arith = lib.GSB_ARITH_NATIVE if (a.arith == "native" and a.precision == "Float32") else lib.GSB_ARITH_REFERENCE
    f = sim.init_fields(s, mcd, a.precision, device=local, arith=arith, global_extent=g)
    if world > 1:
        sim.connect_nccl(f, sim.broadcast_unique_id(rank))

    stream = torch.cuda.ExternalStream(f.compute_stream(), device=local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def barrier():
        f.sync()
        torch.cuda.synchronize()
        if dist:
            dist.barrier()

    sim.iterate(f, s, mcd, nsteps=max(a.warmup, 3))
    barrier()
    launches0 = f.launch_count()
    with ClockSampler(local) as clk:
        e0.record(stream)
        sim.iterate(f, s, mcd, nsteps=a.steps)
        e1.record(stream)
        barrier()
    ms = e0.elapsed_time(e1)
    launches = f.launch_count() - launches0
    # per-phase times of one more step (halo separately, as the north star asks)
    phases = None
    if world > 1:
        f.timers_enable(True)
        sim.iterate(f, s, mcd, nsteps=1)
        phases = f.timers()
        f.timers_enable(False)
    if dist:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = t.item()
        lt = torch.tensor([launches], dtype=torch.int64, device="cuda")
        dist.all_reduce(lt)
        launches = int(lt.item())
    cells = float(a.L) ** 3 * world
    ms_step = ms / a.steps
    glups = cells / ms_step / 1e6

    # ---- e2e: the same metric end to end through the public C ABI with HOST buffers.
    # (1) "e2e": the reference driver's own loop (GrayScott.main, src/GrayScott.jl:81-96, with the
    #     plotgap = 10 of examples/settings-files.toml): the initial u, v are uploaded from pinned host
    #     memory (gsb_set_ghosted), then K iterate! steps with get_fields to pinned host memory every
    #     plotgap steps, as IO.write_step! does (IO.jl:84-85).  All copies are inside the timed region;
    #     the byte counts are per-step averages over the K steps.
    # (2) "e2e_every_step": the strictest reading -- every single step uploads both ghosted fields and
    #     downloads both results (17 GB over PCIe per step at L = 1024); no real driver does this.
    e2e = e2e_strict = raise_skip = None
    if not a.no_e2e:
        T = np.float32 if a.precision == "Float32" else np.float64
        tdt = torch.float32 if T == np.float32 else torch.float64
        n_g, n_i = (a.L + 2) ** 3, a.L ** 3
        # host memory: two ghosted buffers (the upload), reused as the destination of the outputs; the
        # run is skipped rather than risking the box if that does not fit comfortably
        need = 2 * n_g * (4 if T == np.float32 else 8) * int(os.environ.get("LOCAL_WORLD_SIZE", world))
        avail = mem_available_bytes()
        if avail is not None and need > 0.4 * avail:
            raise_skip = f"e2e skipped: {need / 1e9:.0f} GB of pinned host memory needed, {avail / 1e9:.0f} GB available"
        else:
            raise_skip = None
        if dist:                                   # all ranks take the same decision
            ok = torch.tensor([0 if raise_skip else 1], dtype=torch.int32, device="cuda")
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if ok.item() == 0 and raise_skip is None:
                raise_skip = "e2e skipped: another rank lacks host memory"
    if not a.no_e2e and raise_skip is None:
        hu = torch.empty(n_g, dtype=tdt, pin_memory=True)
        hv = torch.empty(n_g, dtype=tdt, pin_memory=True)
        out = [(hu, hv)]

        def refill():
            sim.check(lib.lib().gsb_get_ghosted(f.handle, lib.GSB_U, hu.data_ptr()))
            sim.check(lib.lib().gsb_get_ghosted(f.handle, lib.GSB_V, hv.data_ptr()))

        refill()
        es = hu.element_size()
        plotgap = 10

        def driver_loop(k):
            sim.check(lib.lib().gsb_set_ghosted(f.handle, lib.GSB_U, hu.data_ptr()))
            sim.check(lib.lib().gsb_set_ghosted(f.handle, lib.GSB_V, hv.data_ptr()))
            n_out = 0
            for step in range(1, k + 1):
                sim.check(lib.lib().gsb_step(f.handle, 1))
                if step % plotgap == 0:
                    ou, ov = out[0]
                    sim.check(lib.lib().gsb_get_fields_async(f.handle, ou.data_ptr(), ov.data_ptr()))
                    n_out += 1
            sim.check(lib.lib().gsb_output_wait(f.handle))
            return n_out

        driver_loop(plotgap)                     # warm-up: allocates the staging buffers, touches the host pages
        refill()                                 # the timed run uploads real ghosted fields again
        barrier()
        t0 = time.perf_counter()
        n_out = driver_loop(a.steps)
        barrier()
        dt = time.perf_counter() - t0
        if dist:
            t = torch.tensor([dt], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = t.item()
        e2e = {"value": cells * a.steps / dt / 1e9, "unit": "GLUPS",
               "h2d_bytes_per_step": 2 * n_g * es * world / a.steps, "d2h_bytes_per_step": 2 * n_i * es * world * n_out / a.steps,
               "steps": a.steps, "outputs": n_out, "plotgap": plotgap, "ms_per_step": dt / a.steps * 1e3,
               "what": ("GrayScott.main's loop through the C ABI with host buffers: gsb_set_ghosted(u, v) from pinned host "
                        "once, then per step gsb_step(1) and every plotgap = 10 steps gsb_get_fields_async to pinned host "
                        "(IO.write_step!), gsb_output_wait at the end; all inside the timed region")}

        def one():
            sim.check(lib.lib().gsb_set_ghosted(f.handle, lib.GSB_U, hu.data_ptr()))
            sim.check(lib.lib().gsb_set_ghosted(f.handle, lib.GSB_V, hv.data_ptr()))
            sim.check(lib.lib().gsb_step(f.handle, 1))
            sim.check(lib.lib().gsb_get_fields(f.handle, out[0][0].data_ptr(), out[0][1].data_ptr()))

        refill()
        one()
        refill()
        barrier()
        t0 = time.perf_counter()
        for _ in range(a.e2e_steps):
            one()
        barrier()
        dt = time.perf_counter() - t0
        if dist:
            t = torch.tensor([dt], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = t.item()
        e2e_strict = {"value": cells * a.e2e_steps / dt / 1e9, "unit": "GLUPS",
                      "h2d_bytes_per_step": 2 * n_g * es * world, "d2h_bytes_per_step": 2 * n_i * es * world,
                      "steps": a.e2e_steps, "ms_per_step": dt / a.e2e_steps * 1e3,
                      "what": "per step: gsb_set_ghosted(u), gsb_set_ghosted(v) from pinned host, gsb_step(1), gsb_get_fields to pinned host"}

    halo_bytes = f.halo_bytes()
    kern = [C.c_int(), C.c_int(), C.c_int(), C.c_int()]
    sim.check(lib.lib().gsb_get_kernel(f.handle, *[C.byref(k) for k in kern]))
    f.close()

    if rank != 0:
        if dist:
            dist.destroy_process_group()
        return

    peak, peak_src = peaks()
    bpc = BYTES_PER_CELL[a.precision]
    # the update kernel is the only kernel of a 1-GPU step; its average launch duration is ms_step.
    achieved = float(a.L) ** 3 * bpc / (ms_step * 1e-3) / 1e9
    key = f"L{a.L}_{a.precision}_{a.arith if a.precision == 'Float32' else 'reference'}_noise{a.noise:g}"
    traffic = ncu_traffic(key)
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": float(a.L) ** 3 * bpc,
                "kernel": f"stencil_{'l1' if kern[0].value == 1 else 'tma'}_kernel ty={kern[1].value} stages={kern[2].value} zchunk={kern[3].value}",
                "note": ("per-GPU algorithmic bytes / per-step device time" +
                         ("" if world == 1 else "; at N>1 the step also holds pack/unpack, NCCL and shell launches"))}
    line = {
        "metric": "GLUPS (cell-updates/s)", "value": glups, "unit": "GLUPS", "n_gpus": world, "steps": a.steps,
        "warmup": max(a.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32" if a.precision == "Float32" else "f64",
        "data": "synthetic (reference initial condition, Simulation_CPU.jl:14-65)",
        "config": workload_config(a, world), "clocks": clk.summary(), "gpu_launches": launches,
        "roofline": roofline,
    }
    if e2e:
        line["e2e"] = e2e
        line["e2e_every_step"] = e2e_strict
    elif not a.no_e2e:
        line["e2e"] = {"value": None, "unit": "GLUPS", "h2d_bytes_per_step": None, "d2h_bytes_per_step": None,
                       "what": raise_skip}
    if phases:
        line["halo"] = {"ms_per_step": phases["halo_ms"], "interior_ms": phases["interior_ms"],
                        "shell_ms": phases["shell_ms"], "bytes_out_per_gpu": halo_bytes,
                        "gbs_vs_770_nvlink": halo_bytes / max(phases["halo_ms"], 1e-9) / 1e6}
    if world == 1 and not a.no_cpu_baseline:
        v, Ls, st, thr, secs = cpu_solver_glups(a.L, a.precision, a.noise, budget_s=20.0)
        line["cpu_baseline"] = {
            "value": v, "unit": "GLUPS", "cores": thr, "kind": "port",
            "sample": (f"L={Ls} cube, {st} steps in {secs:.1f} s, same physics and noise draw per cell; C restatement of "
                       f"the reference's Julia CPU-threaded solver (Julia unavailable on this box), {thr} OpenMP threads")}
    emit(line)
    if dist:
        dist.destroy_process_group()


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
