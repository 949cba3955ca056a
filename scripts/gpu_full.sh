#!/bin/bash
# Full GPU regression: every -m gpu test, then the default bench line.
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 2400 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/t_all.log 2>&1
echo "pytest rc=$?" >> gpurun_out/summary.txt; tail -15 gpurun_out/t_all.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench rc=$?" >> gpurun_out/summary.txt
python -c "
import json; d=json.load(open('gpurun_out/bench.json')); print({k:d[k] for k in ('value','ms_per_step','eager_ms_per_step','gpu_launches')}, d['e2e'], d['roofline']['achieved'], d['roofline_bw']['kernel'], d['roofline_bw']['achieved'])"
cat gpurun_out/summary.txt
