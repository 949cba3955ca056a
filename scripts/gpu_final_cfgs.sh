#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 600 python scripts/long_run.py 400 > gpurun_out/long_run.txt 2>&1; echo "long_run rc=$?"; cat gpurun_out/long_run.txt | cut -c1-330
for c in cfg1 cfg4 cfg5; do
  timeout 600 python bench.py --config $c --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$c.json 2> gpurun_out/bench_$c.err; echo "$c rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/bench_$c.json')); print('$c', d['value'], d['unit'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline_bw']['frac'])" || tail -3 gpurun_out/bench_$c.err
done
