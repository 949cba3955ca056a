#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_models_gpu.py -m gpu -q --timeout 600 -k benchmarked > gpurun_out/t_models.log 2>&1
echo "models rc=$?" >> gpurun_out/summary.txt; tail -3 gpurun_out/t_models.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench rc=$?" >> gpurun_out/summary.txt
python -c "
import json; d=json.load(open('gpurun_out/bench.json')); print({k:d[k] for k in ('value','ms_per_step','eager_ms_per_step','gpu_launches')}, d['e2e'], d['roofline']['achieved'])"
for shp in "64 4096 128" "64 256 512"; do
  n=$(echo $shp | tr ' ' '_')
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:plane_cluster -c 2 -f -o gpurun_out/ncu_norm_$n python scripts/bench_norm.py $shp 2 > gpurun_out/ncu_norm_$n.log 2>&1
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"first_|tapgemm_kernel" -c 6 -f -o gpurun_out/ncu_first python scripts/bench_first.py > gpurun_out/ncu_first.log 2>&1
ls -la gpurun_out
cat gpurun_out/summary.txt
