#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_dlayers_gpu.py tests/test_layers_gpu.py tests/test_models_gpu.py -x -q > gpurun_out/t_bits.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/t_bits.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err
python -c "
import json; d=json.load(open('gpurun_out/bench.json')); print({k:d[k] for k in ('value','ms_per_step','eager_ms_per_step')}, d['e2e']['ms_per_step'], d['clocks'], d['roofline'])"
