#!/usr/bin/env python
"""Splits an `ncu --page source --csv` dump of one kernel at its block-wide barriers and prints, per phase: share of the
stall samples, share of the executed instructions and the leading stall reasons; plus the opcode mix of the whole kernel.
  python scripts/ncu_phases.py gpurun_out/<tag>.source.csv"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
idx = {n: i for i, n in enumerate(rows[h])}
R = [r for r in rows[h + 1:] if len(r) == len(rows[h])]
S = lambda r, n: int(r[idx[n]] or 0)
tot = sum(S(r, "# Samples") for r in R)
ti = sum(S(r, "Instructions Executed") for r in R)
print(rows[0][1] if len(rows[0]) > 1 else "", "\nsamples", tot, "static instructions", len(R), "executed", ti)
ops = collections.Counter(); smp = collections.Counter()
for r in R:
    t = r[1].split(); o = t[1] if t[0].startswith("@") else t[0]
    ops[o] += S(r, "Instructions Executed"); smp[o] += S(r, "# Samples")
for o, c in ops.most_common(14):
    print(f"  {o:20s} executed {100 * c / ti:5.1f} %   samples {100 * smp[o] / tot:5.1f} %")
stalls = [n for n in rows[h] if n.startswith("stall_") and "Not Issued" not in n]
bars = [i for i, r in enumerate(R) if "BAR.SYNC" in r[1]]
prev = 0
for b in bars + [len(R) - 1]:
    seg = R[prev:b + 1]
    s = sum(S(r, "# Samples") for r in seg); e = sum(S(r, "Instructions Executed") for r in seg)
    st = sorted(((sum(S(r, k) for r in seg), k[6:]) for k in stalls), reverse=True)[:5]
    print(f"phase [{prev:5d},{b:5d}] samples {100 * s / tot:5.1f} %  executed {100 * e / ti:5.1f} %  ",
          ", ".join(f"{k} {100 * v / max(s, 1):.0f}%" for v, k in st))
    prev = b + 1
