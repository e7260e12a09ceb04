#!/bin/bash
# multi-GPU evidence on N GPUs of one box: two-device tests, --verify, weak + strong bench lines
N=${1:-2}; T=${2:-r2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 python -m pytest tests/test_multi_gpu.py -x -q -m gpu -p no:cacheprovider > gpurun_out/${T}_multi_tests.log 2>&1; echo "multi tests exit $?"; tail -n 2 gpurun_out/${T}_multi_tests.log
timeout 600 $TR bench.py --gpus $N --verify > gpurun_out/${T}_verify_${N}gpu.json 2> gpurun_out/${T}_verify_${N}gpu.err; echo "verify exit $?"; cat gpurun_out/${T}_verify_${N}gpu.json
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/${T}_bench_${N}gpu.json 2> gpurun_out/${T}_bench_${N}gpu.err; echo "weak exit $?"; python tools/show_bench.py gpurun_out/${T}_bench_${N}gpu.json | head -1
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 --scaling strong --batch 256 --no-cpu-baseline > gpurun_out/${T}_bench_${N}gpu_strong.json 2> gpurun_out/${T}_bench_${N}gpu_strong.err; echo "strong exit $?"; python tools/show_bench.py gpurun_out/${T}_bench_${N}gpu_strong.json | head -1
timeout 600 $TR bench.py --gpus $N --steps 10 --warmup 3 --case deit_base_ibert_w16 --batch 64 --no-cpu-baseline > gpurun_out/${T}_bench_${N}gpu_w16.json 2> gpurun_out/${T}_bench_${N}gpu_w16.err; echo "w16 exit $?"; python tools/show_bench.py gpurun_out/${T}_bench_${N}gpu_w16.json | head -1
