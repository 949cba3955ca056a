#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
nvidia-smi -L > gpurun_out/gpus.txt
timeout 900 python -m pytest tests/test_dp2_gpu.py -m gpu -q --timeout 800 -s > gpurun_out/t_dp2.log 2>&1
echo "dp2 rc=$?" >> gpurun_out/summary.txt; tail -25 gpurun_out/t_dp2.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err
echo "bench n2 rc=$?" >> gpurun_out/summary.txt
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
python -c "
import json
for n in (1,2):
    d=json.load(open(f'gpurun_out/bench_n{n}.json')); print(n, {k:d[k] for k in ('value','ms_per_step','eager_ms_per_step')}, d['e2e']['ms_per_step'])"
tail -3 gpurun_out/bench_n2.err
cat gpurun_out/summary.txt
