#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
for v in 0 1 0 1; do
B2C_PDL=$v timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_pdl$v.json 2> gpurun_out/bench_pdl$v.err
python -c "
import json; d=json.load(open('gpurun_out/bench_pdl$v.json')); print('PDL=$v', {k:d[k] for k in ('value','ms_per_step','eager_ms_per_step')}, d['e2e']['ms_per_step'])"
done
