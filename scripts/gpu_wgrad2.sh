#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 300 python -m pytest tests/test_kernels_gpu.py tests/test_simple_gpu.py -x -q -k "wgrad" > gpurun_out/t_wgrad.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/t_wgrad.log
timeout 300 python scripts/bench_gemm.py wgrad 5 > gpurun_out/wgrad2.txt 2>&1; grep -v det gpurun_out/wgrad2.txt
