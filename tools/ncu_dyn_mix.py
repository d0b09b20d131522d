#!/usr/bin/env python
"""Dynamic opcode mix of a kernel from an ncu report's source page.

  ncu -i X.ncu-rep --page source --csv --print-source sass | python tools/ncu_dyn_mix.py <warp_iterations>

Sums "Instructions Executed" (warp-level) per opcode, and prints them per warp-iteration
(warp_iterations = cells / (32 lanes x cells per thread)) plus the stall samples per opcode."""
import collections
import csv
import re
import sys

iters = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
rows = list(csv.reader(sys.stdin))
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hdr]
ci, cs, cx = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
ex, st = collections.Counter(), collections.Counter()
for r in rows[hdr + 1:]:
    if len(r) <= cx:
        continue
    t = re.sub(r"^@!?U?P\d+\s+", "", r[ci].strip())
    op = t.split()[0] if t else "?"
    parts = op.split(".")
    key = parts[0] + ("." + parts[1] if len(parts) > 1 and parts[0] in ("F2F", "I2F", "IMAD", "LDS", "STG", "LDG", "SYNCS", "F2FP", "I2FP") else "")
    ex[key] += int(r[cx] or 0)
    st[key] += int(r[cs] or 0)
tot, tots = sum(ex.values()), sum(st.values())
print(f"total warp instructions {tot}  = {tot / iters:.1f} per warp-iteration; stall samples {tots}")
for k, n in ex.most_common(40):
    print(f"  {k:<14} {n / iters:8.2f} /iter  {100.0 * n / tot:5.1f}%   stall samples {100.0 * st[k] / max(tots, 1):5.1f}%")
