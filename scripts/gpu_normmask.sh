#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_kernels2_gpu.py tests/test_layers_gpu.py tests/test_zz_no_alloc.py -x -q > gpurun_out/t_norm.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/t_norm.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err
python -c "
import json; d=json.load(open('gpurun_out/bench.json')); print({k:d[k] for k in ('value','ms_per_step','eager_ms_per_step')}, d['e2e']['ms_per_step'], d['roofline_bw']['per_kernel']['b2c_in_act_bwd'], d['roofline_bw']['per_kernel']['b2c_in_act_fwd'])"
