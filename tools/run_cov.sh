#!/bin/bash
mkdir -p gpurun_out
T=${1:-cov}
timeout 1500 python -m pytest tests -q -m gpu --timeout 400 -p no:cacheprovider > gpurun_out/${T}_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/${T}_tests.log
grep -v Warning gpurun_out/${T}_tests.log | tail -n 30
