#!/bin/bash
# parity + bench, then an `ncu --set full` capture of one block of the second forward
T=${1:-q}
bash tools/run_quick.sh $T
grep -q "tests exit 0" gpurun_out/${T}_tests.log || exit 1
bash tools/run_ncu_full.sh $T > /dev/null 2>&1
tail -n 3 gpurun_out/${T}_ncu.log
