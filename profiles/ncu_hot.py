#!/usr/bin/env python3
"""Top SASS instructions by stall samples from `ncu --page source --csv --print-source sass` output
(first kernel block only):  ncu_hot.py src.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
N = int(sys.argv[2]) if len(sys.argv) > 2 else 40
starts = [i for i, r in enumerate(rows) if r and r[0] == 'Kernel Name']
end = starts[1] if len(starts) > 1 else len(rows)
hdr = rows[starts[0] + 1]
ia, isrc, isamp, iex = hdr.index('Address'), hdr.index('Source'), hdr.index('# Samples'), hdr.index('Instructions Executed')
stall = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
body = [r for r in rows[starts[0] + 2:end] if len(r) > max(isamp, iex)]
tot = sum(int(r[isamp]) for r in body)
totex = sum(int(r[iex]) for r in body)
print('total samples', tot, 'instructions executed', totex, 'SASS lines', len(body))
agg = {}
for r in body:
    for i in stall:
        agg[hdr[i]] = agg.get(hdr[i], 0) + int(r[i])
print({k: v for k, v in sorted(agg.items(), key=lambda x: -x[1]) if v})
ops = {}
for r in body:
    op = r[isrc].strip().split()
    op = (op[1] if op and op[0].startswith('@') and len(op) > 1 else (op[0] if op else '?')).split('.')[0]
    ops[op] = ops.get(op, 0) + int(r[iex])
print('executed by opcode:', {k: v for k, v in sorted(ops.items(), key=lambda x: -x[1])[:25]})
order = sorted(range(len(body)), key=lambda j: -int(body[j][isamp]))[:N]
for j in sorted(order):
    r = body[j]
    top = sorted(((int(r[i]), hdr[i]) for i in stall), reverse=True)[:2]
    print('%5d %6.2f%% ex=%8s %-60s %s' % (j, 100.0 * int(r[isamp]) / tot, r[iex], r[isrc].strip()[:60], ' '.join('%s=%d' % (n, v) for v, n in top if v)))
