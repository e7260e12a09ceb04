#!/bin/bash
mkdir -p gpurun_out
T=${1:-l}
timeout 200 python -m pytest tests/test_kernels_gpu.py -q --timeout 100 -p no:cacheprovider -x -k layernorm > gpurun_out/${T}_ltests.log 2>&1
echo "tests exit $?" >> gpurun_out/${T}_ltests.log
tail -n 15 gpurun_out/${T}_ltests.log
timeout 100 python tools/ln_bench.py
