#!/usr/bin/env python
"""Opcode mix of the hot loop of one kernel from `cuobjdump -sass` text.

  cuobjdump -sass -fun <mangled> file.o | python tools/sass_mix.py [--iters N]

Finds the longest backward branch (the main loop), prints the opcode histogram of its body and the
total per iteration divided by N (N = unrolled iterations the body holds)."""
import collections
import re
import sys

iters = 1
if "--iters" in sys.argv:
    iters = int(sys.argv[sys.argv.index("--iters") + 1])
ins = []
for ln in sys.stdin:
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
    if m:
        ins.append((int(m.group(1), 16), m.group(2).strip()))
best = None
for i, (addr, text) in enumerate(ins):
    m = re.search(r"\bBRA\b.*?0x([0-9a-f]+)", text)
    if m:
        tgt = int(m.group(1), 16)
        if tgt < addr and (best is None or addr - tgt > best[1] - best[0]):
            best = (tgt, addr)
if not best:
    sys.exit("no backward branch found")
body = [t for a, t in ins if best[0] <= a <= best[1]]
hist = collections.Counter()
for t in body:
    t = re.sub(r"^@!?U?P\d+\s+", "", t)
    op = t.split()[0]
    hist[op.split(".")[0] + ("." + op.split(".")[1] if op.startswith(("LDS", "STG", "LDG", "F2F", "I2F", "IMAD", "UTMALDG", "SYNCS")) and "." in op else "")] += 1
print(f"loop 0x{best[0]:x}..0x{best[1]:x}: {len(body)} instructions, {len(body) / iters:.1f} per iteration (x{iters})")
for op, n in hist.most_common():
    print(f"  {op:<18} {n:5d}  {n / iters:7.1f}")
