#!/bin/bash
# GEMM-only iteration: parity of the GEMM tests (short timeout), then per-shape timing
mkdir -p gpurun_out
T=${1:-g}
timeout 200 python -m pytest tests/test_kernels_gpu.py -q --timeout 100 -p no:cacheprovider -x -k gemm > gpurun_out/${T}_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/${T}_tests.log
tail -n 25 gpurun_out/${T}_tests.log
grep -q "tests exit 0" gpurun_out/${T}_tests.log || exit 1
timeout 200 python tools/gemm_bench.py > gpurun_out/${T}_gemm.log 2>&1
cat gpurun_out/${T}_gemm.log
