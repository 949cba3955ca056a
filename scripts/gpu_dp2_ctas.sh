#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
run() { # name, env...
  name=$1; shift
  env "$@" timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err
  python -c "
import json
d=json.load(open('gpurun_out/bench_$name.json')); print('$name', {k:d[k] for k in ('value','ms_per_step','n_gpus')}, d['e2e']['ms_per_step'], d['clocks']['sm_mhz'])"
}
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
python -c "
import json
d=json.load(open('gpurun_out/bench_n1.json')); print('n1', {k:d[k] for k in ('value','ms_per_step','n_gpus')}, d['e2e']['ms_per_step'], d['clocks']['sm_mhz'])"
run default X=1
run ctas4 NCCL_MAX_CTAS=4
run ctas2 NCCL_MAX_CTAS=2
run ctas1 NCCL_MAX_CTAS=1
run default_b X=1
