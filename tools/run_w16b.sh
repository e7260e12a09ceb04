#!/bin/bash
mkdir -p gpurun_out
T=${1:-w16b}
timeout 900 python -m pytest tests/test_module_api_gpu.py tests/test_plan_gpu.py -q --timeout 400 -p no:cacheprovider > gpurun_out/${T}_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/${T}_tests.log
tail -n 25 gpurun_out/${T}_tests.log
for c in deit_base_ibert_w16 deit_base_ibert deit_tiny_ivit_w16; do
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --case $c --batch 64 > gpurun_out/${T}_bench_$c.json 2> gpurun_out/${T}_bench_$c.err
echo "bench $c exit $?"
tail -n 3 gpurun_out/${T}_bench_$c.err
python - <<PY
import json
d=json.loads(open("gpurun_out/${T}_bench_$c.json").read().strip().splitlines()[-1])
print("$c value", round(d["value"]), "ms", round(d["ms_per_step"],3), "e2e", d["e2e"] and round(d["e2e"]["value"]))
for k,a in d["kernels"].items(): print(f"  {k:20s} {a['launches']:4.0f} {a['ms']:8.3f} ms {a['share']:6.1%} TOP/s={a['TOP/s']} GB/s={a['GB/s']}")
PY
done
