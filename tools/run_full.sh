#!/bin/bash
# full validation: GPU parity suite, smoke, both bench arms, then an ncu launch list of the bench command
mkdir -p gpurun_out
T=${1:-full}
nvidia-smi --query-gpu=name,driver_version,memory.total --format=csv > gpurun_out/${T}_gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider > gpurun_out/${T}_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/${T}_tests.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/${T}_smoke.log 2>&1
echo "smoke exit $?" >> gpurun_out/${T}_smoke.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err
echo "bench ref exit $?" >> gpurun_out/${T}_bench_ref.err
timeout 900 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
echo "bench exit $?" >> gpurun_out/${T}_bench.err
timeout 300 python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/${T}_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${T}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/${T}_ncu.log 2>&1
echo "ncu exit $?" >> gpurun_out/${T}_ncu.log
tail -n 6 gpurun_out/${T}_tests.log gpurun_out/${T}_smoke.log gpurun_out/${T}_bench_ref.err gpurun_out/${T}_bench.err gpurun_out/${T}_ncu.log
cat gpurun_out/${T}_bench_ref.json gpurun_out/${T}_bench.json
