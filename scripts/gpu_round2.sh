#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_kernels2_gpu.py -m gpu -q --timeout 300 -x > gpurun_out/t_kernels.log 2>&1
echo "kernels rc=$?" >> gpurun_out/summary.txt; tail -3 gpurun_out/t_kernels.log
timeout 1500 python -m pytest tests/test_dlayers_gpu.py tests/test_models_gpu.py tests/test_graph_gpu.py tests/test_sizes_gpu.py -m gpu -q --timeout 600 -s > gpurun_out/t_models.log 2>&1
echo "models rc=$?" >> gpurun_out/summary.txt; tail -12 gpurun_out/t_models.log
python scripts/bench_gemm.py wgrad 5 > gpurun_out/wgrad.txt 2>&1; cat gpurun_out/wgrad.txt
for cfg in "B2C_NORM_CLUSTER=0" "B2C_NORM_CLUSTER=1" "B2C_NORM_RS=8" "B2C_NORM_RS=2" "B2C_NORM_CS=4" "B2C_NORM_CS=4 B2C_NORM_RS=16" "B2C_NORM_CS=2 B2C_NORM_RS=4"; do
  for shp in "64 4096 128" "64 1024 256" "64 256 512"; do
    echo "== $cfg $shp" >> gpurun_out/norm.txt
    env $cfg python scripts/bench_norm.py $shp 10 2>&1 | grep "bwd" >> gpurun_out/norm.txt
  done
done
cat gpurun_out/norm.txt
timeout 600 python scripts/kprof.py 64 256 3 > gpurun_out/kprof.txt 2>&1
echo "kprof rc=$?" >> gpurun_out/summary.txt
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench rc=$?" >> gpurun_out/summary.txt
python -c "
import json; d=json.load(open('gpurun_out/bench.json')); print({k:d[k] for k in ('value','ms_per_step','eager_ms_per_step','gpu_launches')}, d['e2e'], d['roofline']['achieved'], d['roofline_bw'])"
cat gpurun_out/summary.txt
