#!/bin/bash
# quick iteration: parity (kernels + plan) then one bench line
mkdir -p gpurun_out
T=${1:-q}
timeout 300 python -m pytest tests/test_kernels_gpu.py tests/test_plan_gpu.py -q --timeout 120 -p no:cacheprovider -x > gpurun_out/${T}_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/${T}_tests.log
grep -q "tests exit 0" gpurun_out/${T}_tests.log || { tail -n 30 gpurun_out/${T}_tests.log; exit 1; }
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
echo "bench exit $?" >> gpurun_out/${T}_bench.err
tail -n 12 gpurun_out/${T}_tests.log gpurun_out/${T}_bench.err
python - <<PY
import json
d=json.loads(open("gpurun_out/${T}_bench.json").read().strip().splitlines()[-1])
print("value", round(d["value"]), "ms", round(d["ms_per_step"],3), "e2e", d["e2e"] and round(d["e2e"]["value"]), d["e2e"] and d["e2e"]["chunks"])
for k,a in d["kernels"].items(): print(f"  {k:20s} {a['launches']:4.0f} {a['ms']:8.3f} ms {a['share']:6.1%} TOP/s={a['TOP/s']} GB/s={a['GB/s']}")
PY
