#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_kernels_gpu.py -x -q -k "head" > gpurun_out/t_np2.log 2>&1; echo "rc=$?"; tail -6 gpurun_out/t_np2.log | cut -c1-300
timeout 900 python -m pytest tests/test_sizes_gpu.py -x -q > gpurun_out/t_sizes.log 2>&1; echo "rc=$?"; grep -E "Error|error|assert|passed|failed" gpurun_out/t_sizes.log | tail -8 | cut -c1-300
