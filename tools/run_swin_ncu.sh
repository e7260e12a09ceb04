#!/bin/bash
# Swin caller GPU test, then one `ncu --set full` capture of the first block of the second Swin-T forward
mkdir -p gpurun_out
T=${1:-swinn}
timeout 600 python -m pytest tests/test_swin_gpu.py -q --timeout 400 -p no:cacheprovider -x > gpurun_out/${T}_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/${T}_tests.log
tail -n 6 gpurun_out/${T}_tests.log
timeout 300 python tools/prof_swin.py 64 > gpurun_out/${T}_plain.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on \
  -k regex:"window_attention|layernorm_x16|gemm_i8_cl_kernel|patch_merge" -s 60 -c 10 -f -o gpurun_out/${T}_prof \
  python tools/prof_swin.py 64 > gpurun_out/${T}_ncu.log 2>&1
echo "ncu exit $?" >> gpurun_out/${T}_ncu.log
tail -n 3 gpurun_out/${T}_plain.log gpurun_out/${T}_ncu.log
ls -la gpurun_out/ | grep ${T}
