#!/bin/bash
# 8-GPU scaling check: N=8 and N=1 bench lines on the same box (+ cfg4 at N=8)
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
nvidia-smi -L > gpurun_out/gpus.txt
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err
echo "bench n8 rc=$?"
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 8 --steps 10 --warmup 3 --config cfg4 > gpurun_out/bench_cfg4_n8.json 2> gpurun_out/bench_cfg4_n8.err
echo "bench cfg4 n8 rc=$?"
python -c "
import json
for f in ('bench_n1','bench_n8','bench_cfg4_n8'):
    try:
        d=json.load(open(f'gpurun_out/{f}.json')); print(f, {k:d[k] for k in ('value','ms_per_step','n_gpus')}, d['e2e']['value'], d['e2e']['ms_per_step'])
    except Exception as e: print(f, 'FAILED', e)
"
tail -3 gpurun_out/bench_n8.err
