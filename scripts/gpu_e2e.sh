#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_graph_gpu.py tests/test_curve_gpu.py -x -q -s > gpurun_out/t_graph.log 2>&1; echo "rc=$?"; grep -E "deviation|passed|failed|Error" gpurun_out/t_graph.log | tail -8
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err
python -c "
import json; d=json.load(open('gpurun_out/bench.json')); print({k:d[k] for k in ('value','ms_per_step','eager_ms_per_step')}, d['e2e'])"
