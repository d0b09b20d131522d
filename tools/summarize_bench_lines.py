#!/usr/bin/env python
"""One line per bench.py JSON line: what BASELINE.md quotes.   python tools/summarize_bench_lines.py file.json[l] ..."""
import json
import sys

for path in sys.argv[1:]:
    for ln in open(path):
        ln = ln.strip()
        if not ln.startswith("{"):
            continue
        d = json.loads(ln)
        if d.get("impl") == "reference":
            print(f"{path}: CPU arm {d['value']:.3f} GLUPS, {d['cpu_baseline']['cores']} cores, {d['ms_per_step']:.1f} ms/step")
            continue
        w = d.get("windows", {})
        print(f"{path}: N={d['n_gpus']} {d['config']['per_gpu']} {d['config']['precision']} {d['scaling']}: {d['ms_per_step']:.4f} ms "
              f"{d['value']:.1f} GLUPS frac/gpu {d['roofline']['frac']:.3f} windows {[round(x, 3) for x in w.get('ms_per_step', [])]} "
              f"clocks {d['clocks']['reasons']} later {(w.get('clocks_later_windows') or {}).get('reasons')} launches {d['gpu_launches']}")
        if "halo" in d:
            h = d["halo"]
            print(f"    halo pack {h['pack_ms']:.3f} wire {h['wire_ms']:.3f} unpack {h['unpack_ms']:.3f} ms, {h['bytes_out_per_gpu']} B out, "
                  f"pack+unpack max over ranks {h['pack_unpack_ms_max_over_ranks']:.3f}")
        if "parity" in d:
            print(f"    parity ok={d['parity']['ok']} ({len(d['parity']['cases'])} cases)")
        if "strong" in d and "value" in d["strong"]:
            s = d["strong"]
            print(f"    strong {s['config']}: {s['ms_per_step']:.4f} ms {s['value']:.1f} GLUPS frac/gpu {s['roofline_frac_per_gpu']:.3f}")
        if "value_reference_arith" in d:
            r = d["value_reference_arith"]
            print(f"    reference arithmetic {r['ms_per_step']:.3f} ms {r['value']:.1f} GLUPS frac {r['roofline_frac']:.3f}")
        for k in ("e2e", "e2e_async", "e2e_every_step"):
            if k in d and d[k].get("value"):
                print(f"    {k} {d[k]['value']:.2f} GLUPS, PCIe {d[k].get('pcie_gbs_per_gpu', 0):.1f} GB/s per GPU")
        if "cpu_baseline" in d:
            print(f"    cpu_baseline {d['cpu_baseline']['value']:.3f} GLUPS on {d['cpu_baseline']['cores']} cores")
