#!/bin/bash
mkdir -p gpurun_out
T=${1:-swin}
timeout 900 python -m pytest tests/test_swin_gpu.py -q --timeout 400 -p no:cacheprovider -x ${2:-} > gpurun_out/${T}_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/${T}_tests.log
tail -n 60 gpurun_out/${T}_tests.log
