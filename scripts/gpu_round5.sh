#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_kernels2_gpu.py -m gpu -q --timeout 120 -x -k "instance_norm" > gpurun_out/t_kernels.log 2>&1
echo "kernels rc=$?" >> gpurun_out/summary.txt; tail -3 gpurun_out/t_kernels.log
for cfg in "B2C_NORM_PERSIST=0" "B2C_NORM_PERSIST=1" "B2C_NORM_PERSIST=1 B2C_NORM_RS=8"; do
for shp in "64 4096 128" "64 1024 256" "64 256 512" "16 4096 256"; do
  echo "== $cfg" >> gpurun_out/norm.txt
  env $cfg timeout 120 python scripts/bench_norm.py $shp 10 2>&1 | grep "bwd" >> gpurun_out/norm.txt
done
done
cat gpurun_out/norm.txt
cat gpurun_out/summary.txt
