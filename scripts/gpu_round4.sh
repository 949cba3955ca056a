#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_kernels2_gpu.py -m gpu -q --timeout 300 -k "instance_norm" > gpurun_out/t_kernels.log 2>&1
echo "kernels rc=$?" >> gpurun_out/summary.txt; tail -3 gpurun_out/t_kernels.log
for cfg in "B2C_NORM_CLUSTER=1" "B2C_NORM_RS=2" "B2C_NORM_RS=8"; do
  for shp in "64 4096 128" "64 1024 256" "64 256 512"; do
    echo "== $cfg $shp" >> gpurun_out/norm.txt
    env $cfg python scripts/bench_norm.py $shp 10 2>&1 | grep "bwd" >> gpurun_out/norm.txt
  done
done
cat gpurun_out/norm.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:plane_cluster -c 1 -f -o gpurun_out/ncu_norm2 python scripts/bench_norm.py 64 4096 128 2 > gpurun_out/ncu_norm2.log 2>&1
cat gpurun_out/summary.txt
