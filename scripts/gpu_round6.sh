#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_inference_gpu.py tests/test_graph_gpu.py -m gpu -q --timeout 600 > gpurun_out/t_inf.log 2>&1
echo "inference tests rc=$?" >> gpurun_out/summary.txt; tail -12 gpurun_out/t_inf.log
timeout 900 python bench.py --config cfg4 --steps 10 --warmup 3 > gpurun_out/bench_cfg4.json 2> gpurun_out/bench_cfg4.err
echo "cfg4 rc=$?" >> gpurun_out/summary.txt
timeout 900 python bench.py --config cfg5 --steps 10 --warmup 3 > gpurun_out/bench_cfg5.json 2> gpurun_out/bench_cfg5.err
echo "cfg5 rc=$?" >> gpurun_out/summary.txt
timeout 900 python bench.py --config cfg1 --steps 10 --warmup 3 > gpurun_out/bench_cfg1.json 2> gpurun_out/bench_cfg1.err
echo "cfg1 rc=$?" >> gpurun_out/summary.txt
python - <<'PY'
import json
for c in ("cfg4","cfg5","cfg1"):
    try:
        d=json.load(open(f"gpurun_out/bench_{c}.json"))
        print(c, {k:d.get(k) for k in ("value","ms_per_step","eager_ms_per_step")}, d["e2e"], (d.get("roofline") or {}).get("achieved"), (d.get("roofline_bw") or {}).get("kernel"), (d.get("roofline_bw") or {}).get("achieved"), d.get("cpu_baseline"))
    except Exception as e:
        print(c, "failed", e)
PY
tail -3 gpurun_out/bench_cfg4.err gpurun_out/bench_cfg5.err
cat gpurun_out/summary.txt
