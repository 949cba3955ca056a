#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -k "conv" > gpurun_out/t_conv.log 2>&1; echo "rc=$?"; tail -2 gpurun_out/t_conv.log
{ timeout 300 python scripts/bench_gemm.py conv 5; timeout 300 python scripts/bench_gemm.py convt 5 | grep -v "+"; } > gpurun_out/conv.txt 2>&1; cat gpurun_out/conv.txt
