#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_kernels2_gpu.py -m gpu -q --timeout 300 -k "instance_norm" > gpurun_out/t_kernels.log 2>&1
echo "kernels rc=$?" >> gpurun_out/summary.txt; tail -3 gpurun_out/t_kernels.log
for shp in "64 4096 128" "64 1024 256" "64 256 512" "16 16384 128" "16 4096 256" "16 1024 512"; do
  python scripts/bench_norm.py $shp 10 2>&1 | grep "bwd\|fwd" >> gpurun_out/norm.txt
done
cat gpurun_out/norm.txt
cat gpurun_out/summary.txt
