#!/bin/bash
# round-1 call A: kernel + whole-forward parity, smoke, first bench line, ncu launch list
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,memory.total --format=csv > gpurun_out/a_gpu.txt 2>&1
timeout 900 python -m pytest tests/test_kernels_gpu.py -q --timeout 300 -p no:cacheprovider > gpurun_out/a_kernels.log 2>&1
echo "kernels exit $?" >> gpurun_out/a_kernels.log
timeout 900 python -m pytest tests/test_plan_gpu.py -q --timeout 600 -p no:cacheprovider > gpurun_out/a_plan.log 2>&1
echo "plan exit $?" >> gpurun_out/a_plan.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/a_smoke.log 2>&1
echo "smoke exit $?" >> gpurun_out/a_smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err
echo "bench exit $?" >> gpurun_out/a_bench.err
timeout 300 python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/a_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/a_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/a_ncu.log 2>&1
echo "ncu exit $?" >> gpurun_out/a_ncu.log
tail -n 15 gpurun_out/a_kernels.log gpurun_out/a_plan.log gpurun_out/a_smoke.log gpurun_out/a_bench.err
cat gpurun_out/a_bench.json
