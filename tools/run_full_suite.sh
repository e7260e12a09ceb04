#!/bin/bash
mkdir -p gpurun_out
T=${1:-full}
timeout 1500 python -m pytest tests -q -m gpu --timeout 400 -p no:cacheprovider > gpurun_out/${T}_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/${T}_tests.log
tail -n 6 gpurun_out/${T}_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke exit $?"; tail -n 2 gpurun_out/${T}_smoke.log
for cb in "deit_base_ibert_w16 64" "deit_small_ivit 256"; do
set -- $cb
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --case $1 --batch $2 > gpurun_out/${T}_bench_$1.json 2> gpurun_out/${T}_bench_$1.err
echo "bench $1 $2 exit $?"
python - <<PY
import json
d=json.loads(open("gpurun_out/${T}_bench_$1.json").read().strip().splitlines()[-1])
print("$1 $2 value", round(d["value"]), "ms", round(d["ms_per_step"],3), "e2e", d["e2e"] and round(d["e2e"]["value"]))
for k,a in d["kernels"].items(): print(f"  {k:20s} {a['launches']:4.0f} {a['ms']:8.3f} ms {a['share']:6.1%} TOP/s={a['TOP/s']} GB/s={a['GB/s']}")
PY
done
