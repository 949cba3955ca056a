#!/bin/bash
# 8-GPU scaling check: N=8 (overlapped slices, then ONE all-reduce per network) and N=1 bench lines on the same box
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
nvidia-smi -L > gpurun_out/gpus.txt
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err
echo "bench n8 rc=$?"
B2C_DP_OVERLAP=0 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n8_noov.json 2> gpurun_out/bench_n8_noov.err
echo "bench n8 no-overlap rc=$?"
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
python -c "
import json
for f in ('bench_n1','bench_n8','bench_n8_noov'):
    try:
        d=json.load(open(f'gpurun_out/{f}.json')); print(f, {k:d[k] for k in ('value','ms_per_step','n_gpus')}, d['e2e']['value'], d['e2e']['ms_per_step'], d['clocks'])
    except Exception as e: print(f, 'FAILED', e)
"
tail -3 gpurun_out/bench_n8.err
