#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel.
usage: summarize_launches.py launches.csv [bench.json] > profiles/<name>.md"""
import collections
import csv
import json
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi, gi, bi = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size"), hdr.index("Block Size")
agg = collections.OrderedDict()
for r in rows[1:]:
    try:
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    n = r[ki].split("(")[0].replace("void ", "")
    a = agg.setdefault(n, [0, 0.0, r[gi], r[bi]])
    a[0] += 1
    a[1] += v
tot = sum(v[1] for v in agg.values())
print(f"# ncu launch list summary: `{sys.argv[1]}`\n")
print("ncu `--metrics gpu__time_duration.sum --clock-control none` (cold-cache, serialised launches: compare SHARES).\n")
print("| kernel | launches | total ms | share | avg us | grid (first) | block |")
print("|---|---|---|---|---|---|---|")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"| `{k}` | {v[0]} | {v[1] / 1e6:.3f} | {v[1] / tot:.1%} | {v[1] / v[0] / 1e3:.1f} | {v[2]} | {v[3]} |")
if len(sys.argv) > 2:
    d = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print("\n## bench.py line of the same build (CUDA events, not under ncu)\n")
    e2e = d.get("e2e") or {}
    print(f"- value: {d['value']:.0f} {d['unit']} ({d['ms_per_step']:.3f} ms/step, batch "
          f"{d['config']['batch_per_gpu']}); e2e: {e2e.get('value')}")
    print(f"- clocks: {d.get('clocks')}")
    print("\n| kernel (events) | launches/step | ms/step | share | TOP/s | GB/s |\n|---|---|---|---|---|---|")
    for k, a in d.get("kernels", {}).items():
        print(f"| {k} | {a['launches']:.0f} | {a['ms']} | {a['share']:.1%} | {a['TOP/s']} | {a['GB/s']} |")
    print(f"\nroofline: `{json.dumps(d.get('roofline'))}`")
