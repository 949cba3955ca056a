#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_refine_gpu.py -m gpu -q --timeout 600 > gpurun_out/t_refine.log 2>&1
echo "refine tests rc=$?" >> gpurun_out/summary.txt; tail -25 gpurun_out/t_refine.log
cat gpurun_out/summary.txt
