#!/bin/bash
mkdir -p gpurun_out
T=${1:-w16c}
timeout 900 python -m pytest tests/test_kernels16_gpu.py tests/test_swin_gpu.py tests/test_plan_gpu.py -q --timeout 400 -p no:cacheprovider -x > gpurun_out/${T}_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/${T}_tests.log
tail -n 25 gpurun_out/${T}_tests.log
grep -q "tests exit 0" gpurun_out/${T}_tests.log || exit 1
for cb in "swin_tiny 256" "deit_base_ibert_w16 64" "deit_small_ivit 256"; do
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
