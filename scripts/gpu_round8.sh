#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 300 python -m pytest tests/test_kernels_gpu.py tests/test_simple_gpu.py -m gpu -q --timeout 120 -x -k "conv" > gpurun_out/t_k.log 2>&1
echo "kernel tests rc=$?" >> gpurun_out/summary.txt; tail -3 gpurun_out/t_k.log
for v in 1; do
  echo "== B2C_TAPGEMM_2CTA=$v" >> gpurun_out/gemm.txt
  B2C_TAPGEMM_2CTA=$v timeout 300 python scripts/bench_gemm.py conv 5 >> gpurun_out/gemm.txt 2>&1
  B2C_TAPGEMM_2CTA=$v timeout 300 python scripts/bench_gemm.py convt 5 >> gpurun_out/gemm.txt 2>&1
done
cat gpurun_out/gemm.txt
