#!/bin/bash
# Where do the +0.4 ms at N=2 come from: the collective, or two GPUs that do not run alike?
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
one() { CUDA_VISIBLE_DEVICES=$1 timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/w_$2.json 2>/dev/null; }
show() { python -c "
import json; d=json.load(open('gpurun_out/w_$1.json')); print('$1', d['ms_per_step'], d['clocks'])"; }
one 0 gpu0_alone; show gpu0_alone
one 1 gpu1_alone; show gpu1_alone
one 0 gpu0_both & one 1 gpu1_both & wait; show gpu0_both; show gpu1_both
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/w_n2.json 2>/dev/null; show n2
B2C_DP_OVERLAP=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/w_n2_noov.json 2>/dev/null; show n2_noov
nvidia-smi --query-gpu=index,power.limit,clocks.max.sm,temperature.gpu --format=csv
