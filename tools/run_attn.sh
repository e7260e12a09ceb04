#!/bin/bash
# attention-only iteration: parity of the attention tests (short timeout), then the quick script
mkdir -p gpurun_out
T=${1:-a}
timeout 200 python -m pytest tests/test_kernels_gpu.py -q --timeout 100 -p no:cacheprovider -x -k attention > gpurun_out/${T}_atests.log 2>&1
echo "tests exit $?" >> gpurun_out/${T}_atests.log
tail -n 15 gpurun_out/${T}_atests.log
grep -q "tests exit 0" gpurun_out/${T}_atests.log || exit 1
bash tools/run_quick.sh $T
