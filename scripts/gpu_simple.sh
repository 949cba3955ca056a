#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_simple_gpu.py tests/test_dataset_gpu.py -x -q -s > gpurun_out/t_simple.log 2>&1; tail -25 gpurun_out/t_simple.log
