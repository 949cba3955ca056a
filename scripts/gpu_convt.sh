#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -k "convt or conv4s2 or fp32_check or fused_lrelu" > gpurun_out/t_convt.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/t_convt.log
timeout 300 python scripts/bench_gemm.py convt 5 > gpurun_out/convt.txt 2>&1; cat gpurun_out/convt.txt
