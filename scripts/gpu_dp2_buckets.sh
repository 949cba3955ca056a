#!/bin/bash
# N=2: how the gradient all-reduce is cut / overlapped (B2C_DP_BUCKET_MB, B2C_DP_OVERLAP) and NCCL's CTA count.
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
run() {  # name, env...
  name=$1; shift
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/dp2_$name.json 2> gpurun_out/dp2_$name.err
  python -c "
import json; d=json.load(open('gpurun_out/dp2_$name.json')); print('$name', d['value'], d['ms_per_step'], d['e2e']['ms_per_step'])" || tail -3 gpurun_out/dp2_$name.err
}
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/dp2_n1.json 2>/dev/null; python -c "
import json; d=json.load(open('gpurun_out/dp2_n1.json')); print('n1', d['value'], d['ms_per_step'])"
run default X=1
run b16 B2C_DP_BUCKET_MB=16
run b64 B2C_DP_BUCKET_MB=64
run nooverlap B2C_DP_OVERLAP=0
run minctas32 NCCL_MIN_CTAS=32
run b64_minctas32 B2C_DP_BUCKET_MB=64 NCCL_MIN_CTAS=32
run default2 X=1
