#!/usr/bin/env python
"""Digest of `ncu --page source --csv`: executed warp-instructions and stall samples per phase (split at
block barriers / MMA commits), plus an opcode histogram.  First kernel of the file only."""
import csv
import sys
from collections import Counter

rows = list(csv.reader(open(sys.argv[1])))
h = rows[1]
isrc, iex, ist = h.index('Source'), h.index('Instructions Executed'), h.index('Warp Stall Sampling (All Samples)')
data = []
for r in rows[2:]:
    if len(r) <= iex or not r[0].startswith('0x'):
        if r and r[0] == 'Kernel Name':
            break
        continue
    data.append((int(r[iex] or 0), int(r[ist] or 0), r[isrc].strip()))
tot = sum(d[0] for d in data)
tst = sum(d[1] for d in data)
print(f"warp-instructions executed {tot}, stall samples {tst}, SASS lines {len(data)}")
acc_ex = acc_st = 0
start = 0
for i, (ex, st, s) in enumerate(data):
    acc_ex += ex
    acc_st += st
    if 'BAR.SYNC' in s or 'EXIT' in s or 'UTCBAR' in s:
        print(f"{start:5d}-{i:5d} ex={acc_ex:10d} ({acc_ex / tot:6.1%}) stall={acc_st:7d} ({acc_st / max(tst, 1):6.1%}) | {s[:60]}")
        acc_ex = acc_st = 0
        start = i + 1
c, cs = Counter(), Counter()
for ex, st, s in data:
    t = s.split()
    op = t[1] if s.startswith('@') and len(t) > 1 else t[0]
    op = op.split('.')[0]
    c[op] += ex
    cs[op] += st
for op, n in c.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 30):
    print(f"{op:12s} {n:10d} {n / tot:6.1%}  stall {cs[op] / max(tst, 1):6.1%}")
