#!/bin/bash
# 16-bit locations: kernel + plan parity first, then the whole GPU suite, then one bench line
mkdir -p gpurun_out
T=${1:-w16}
timeout 600 python -m pytest tests/test_kernels16_gpu.py tests/test_plan_gpu.py -q --timeout 300 -p no:cacheprovider -k "16 or mixed_bw or softmax16" > gpurun_out/${T}_tests16.log 2>&1
echo "tests16 exit $?" >> gpurun_out/${T}_tests16.log
timeout 600 python -m pytest tests/test_kernels_gpu.py -q --timeout 300 -p no:cacheprovider -k "softmax16" >> gpurun_out/${T}_tests16.log 2>&1
echo "softmax16 exit $?" >> gpurun_out/${T}_tests16.log
tail -n 60 gpurun_out/${T}_tests16.log
timeout 900 python -m pytest tests -q -m gpu --timeout 300 -p no:cacheprovider > gpurun_out/${T}_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/${T}_tests.log
tail -n 15 gpurun_out/${T}_tests.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
echo "bench exit $?" >> gpurun_out/${T}_bench.err
tail -n 5 gpurun_out/${T}_bench.err
python - <<PY
import json
d=json.loads(open("gpurun_out/${T}_bench.json").read().strip().splitlines()[-1])
print("value", round(d["value"]), "ms", round(d["ms_per_step"],3), "e2e", d["e2e"] and round(d["e2e"]["value"]))
PY
