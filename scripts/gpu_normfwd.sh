#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_kernels2_gpu.py tests/test_layers_gpu.py -x -q -k "norm or layer" > gpurun_out/t_norm.log 2>&1; tail -3 gpurun_out/t_norm.log
{
for shape in "64 256 512" "64 1024 256" "64 4096 128" "64 16384 64" "16 16384 128"; do
  for f in 0 1; do echo "== $shape flat=$f"; B2C_NORM_FLAT=$f timeout 120 python scripts/bench_norm.py $shape 10 2>&1 | grep -E "in_act_fwd"; done
done
} > gpurun_out/norm_fwd.txt 2>&1
cat gpurun_out/norm_fwd.txt
