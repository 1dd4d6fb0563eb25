#!/usr/bin/env python3
"""Summarise an `ncu --page source --csv` export: shared-memory wavefronts (actual / ideal / excessive) and executed
instructions per opcode class, plus the instructions with the most excessive wavefronts.   ncu_source_summary.py file.csv"""
import csv, sys, collections, re
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
def f(r, name):
    try: return float(r[ix[name]])
    except Exception: return 0.0
by = collections.defaultdict(lambda: [0, 0, 0, 0, 0])
tot = [0, 0, 0, 0, 0]
worst = []
for r in rows[2:]:
    if len(r) < len(hdr): continue
    src = r[ix["Source"]].strip()
    m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)", src)
    op = m.group(2) if m else "?"
    key = op.split(".")[0] + ("." + op.split(".")[1] if op.startswith(("LDS", "STS", "LDG", "STG", "LDGSTS")) and "." in op else "")
    v = [f(r, "Instructions Executed"), f(r, "L1 Wavefronts Shared"), f(r, "L1 Wavefronts Shared Ideal"), f(r, "L1 Wavefronts Shared Excessive"), f(r, "# Samples")]
    for q in range(5): by[key][q] += v[q]; tot[q] += v[q]
    if v[3] > 0: worst.append((v[3], v[1], v[2], v[0], src))
print("total: inst %.0f  smem wavefronts %.0f ideal %.0f excessive %.0f  samples %.0f" % tuple(tot))
print("%-14s %12s %12s %12s %12s %8s" % ("opcode", "inst", "wavefronts", "ideal", "excessive", "samples"))
for k, v in sorted(by.items(), key=lambda kv: -kv[1][0])[:28]:
    print("%-14s %12.0f %12.0f %12.0f %12.0f %8.0f" % (k, *v))
worst.sort(reverse=True)
print("instructions with the most excessive shared wavefronts:")
for w in worst[:int(sys.argv[2]) if len(sys.argv) > 2 else 25]:
    print("  excess %9.0f  wavefronts %9.0f ideal %9.0f inst %8.0f  %s" % w)
