#!/usr/bin/env python
"""Digest of `ncu -i X.ncu-rep --page source --csv`: executed instructions and stall samples per
region between marker instructions (LDTM / SYNCS / STG / UTCIMMA / UTMALDG ...), first kernel only."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
h = rows[1]
isrc, iex, ist = h.index('Source'), h.index('Instructions Executed'), h.index('Warp Stall Sampling (All Samples)')
data = []
for r in rows[2:]:
    if len(r) <= iex or r[0] == 'Address' or not r[0].startswith('0x'):
        if r and r[0] == 'Kernel Name':
            break
        continue
    data.append((int(r[iex] or 0), int(r[ist] or 0), r[isrc].strip()))
tot = sum(d[0] for d in data)
tst = sum(d[1] for d in data)
print(f"instructions executed {tot}, stall samples {tst}, SASS lines {len(data)}")
marks = ('LDTM', 'SYNCS', 'UTCIMMA', 'UTMALDG', 'UTCBAR', 'EXIT', 'BAR.', 'WARPSYNC.ALL')
acc_ex = acc_st = 0
start = 0
for i, (ex, st, s) in enumerate(data):
    acc_ex += ex
    acc_st += st
    if any(m in s for m in marks):
        print(f"{start:5d}-{i:5d} ex={acc_ex:10d} ({acc_ex / tot:6.1%}) stall={acc_st:7d} ({acc_st / max(tst, 1):6.1%})  | {ex:8d} {s[:70]}")
        acc_ex = acc_st = 0
        start = i + 1
if len(sys.argv) > 2:
    lo, hi = int(sys.argv[2]), int(sys.argv[3])
    for i in range(lo, hi):
        print(f"{i:5d} {data[i][0]:9d} {data[i][1]:6d} {data[i][2][:90]}")
