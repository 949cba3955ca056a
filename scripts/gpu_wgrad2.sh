#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 300 python -m pytest tests/test_kernels_gpu.py -x -q -k "wgrad4s2" > gpurun_out/t_wgrad.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/t_wgrad.log
{ echo "== 1-CTA"; B2C_WGRAD_2CTA=0 timeout 300 python scripts/bench_gemm.py wgrad 5; echo "== pairs"; timeout 300 python scripts/bench_gemm.py wgrad 5; } > gpurun_out/wgrad2.txt 2>&1; cat gpurun_out/wgrad2.txt
