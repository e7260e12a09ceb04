#!/bin/bash
mkdir -p gpurun_out
T=${1:-winn}
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"window_attention" -s 12 -c 2 -f -o gpurun_out/${T}_prof \
  python tools/prof_swin.py 64 > gpurun_out/${T}_ncu.log 2>&1
echo "ncu exit $?"; tail -n 2 gpurun_out/${T}_ncu.log
