#!/usr/bin/env python
"""Opcode histogram of an ncu SASS source page (ncu -i rep --page source --csv --print-source sass), weighted by
executed warp instructions; optional second argument = samples processed, to print instructions per sample."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
iS, iE, iW = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
h, st = collections.Counter(), collections.Counter()
tot = 0
for r in rows[2:]:
    if len(r) <= iE: continue
    src = r[iS].strip()
    if src.startswith('@'): src = src.split(None, 1)[1]
    op = src.split()[0].split('.')[0]
    n = int(r[iE] or 0); tot += n
    h[op] += n; st[op] += int(r[iW] or 0)
samples = float(sys.argv[2]) if len(sys.argv) > 2 else None
print("total warp instr", tot, "" if not samples else f"= {tot*32/samples:.2f} thread-instr/sample")
for op, n in h.most_common(40):
    extra = f"{n*32/samples:7.3f}/sample" if samples else ""
    print(f"{op:12s} {n:12d} {100.0*n/tot:6.2f}%  stall-samples {st[op]:7d} {extra}")
