"""Pretty-print the JSON line bench.py wrote (last line of the file)."""
import json
import sys

d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
e = d.get("e2e")
print("value", round(d["value"]), "ms", round(d["ms_per_step"], 3), "e2e", e and round(e["value"]),
      "cpu", (d.get("cpu_baseline") or {}).get("value"), "roof", (d.get("roofline") or {}).get("frac"), d.get("clocks"))
for k, a in (d.get("kernels") or {}).items():
    print(f"  {k:22s} {a['launches']:4.0f} {a['ms']:8.3f} ms {a['share']:6.1%} TOP/s={a.get('TOP/s')} GB/s={a.get('GB/s')}")
