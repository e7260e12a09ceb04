#!/bin/bash
# round-end evidence: whole GPU suite, smoke, default bench line (with cpu_baseline), reference arm, ncu launch list
mkdir -p gpurun_out
T=${1:-final}
timeout 1500 python -m pytest tests -q -m gpu --timeout 400 -p no:cacheprovider > gpurun_out/${T}_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/${T}_tests.log
grep -v -i warning gpurun_out/${T}_tests.log | tail -n 4
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke exit $?"; tail -n 1 gpurun_out/${T}_smoke.log
timeout 600 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench exit $?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref exit $?"
python - <<PY
import json
d=json.loads(open("gpurun_out/${T}_bench.json").read().strip().splitlines()[-1])
print("value", round(d["value"]), "ms", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"]), "cpu", d["cpu_baseline"]["value"], "roof", d["roofline"]["frac"], d["clocks"])
r=json.loads(open("gpurun_out/${T}_bench_ref.json").read().strip().splitlines()[-1]); print("ref", r["value"], r["cpu_baseline"]["cores"])
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/${T}_ncu.log 2>&1
echo "ncu exit $?"
