#!/bin/bash
# first bring-up: kernel-level parity, three isolated processes (a fault in one must not poison the others)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,memory.total --format=csv > gpurun_out/k1_gpu.txt 2>&1
timeout 600 python -m pytest tests/test_kernels_gpu.py -q --timeout 300 -p no:cacheprovider -k "not gemm and not attention" > gpurun_out/k1_row.log 2>&1
echo "row exit $?" >> gpurun_out/k1_row.log
timeout 600 python -m pytest tests/test_kernels_gpu.py -q --timeout 300 -p no:cacheprovider -k "gemm" > gpurun_out/k1_gemm.log 2>&1
echo "gemm exit $?" >> gpurun_out/k1_gemm.log
timeout 600 python -m pytest tests/test_kernels_gpu.py -q --timeout 300 -p no:cacheprovider -k "attention" > gpurun_out/k1_attn.log 2>&1
echo "attn exit $?" >> gpurun_out/k1_attn.log
tail -5 gpurun_out/k1_row.log gpurun_out/k1_gemm.log gpurun_out/k1_attn.log
