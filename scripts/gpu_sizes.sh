#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_kernels_gpu.py -x -q -k "conv4s2 or convt or wgrad4s2" > gpurun_out/t_np2.log 2>&1; echo "rc=$?"; tail -6 gpurun_out/t_np2.log | cut -c1-300
timeout 900 python -m pytest tests/test_sizes_gpu.py -x -q > gpurun_out/t_sizes.log 2>&1; echo "rc=$?"; tail -8 gpurun_out/t_sizes.log | cut -c1-300
