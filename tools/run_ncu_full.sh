#!/bin/bash
# one `ncu --set full` capture of the second forward's first block (patch GEMM, LN, qkv, attention, proj, LN, fc1, fc2)
mkdir -p gpurun_out
T=${1:-ncu}
timeout 300 python tools/prof_step.py > gpurun_out/${T}_plain.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on \
  -k regex:"layernorm_ivit|gemm_i8_|attention_i8|gelu_lut_apply" -s 99 -c 9 -f -o gpurun_out/${T}_prof \
  python tools/prof_step.py > gpurun_out/${T}_ncu.log 2>&1
echo "ncu exit $?" >> gpurun_out/${T}_ncu.log
tail -n 5 gpurun_out/${T}_plain.log gpurun_out/${T}_ncu.log
ls -la gpurun_out/
