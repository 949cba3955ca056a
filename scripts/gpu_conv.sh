#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_dlayers_gpu.py tests/test_layers_gpu.py -x -q > gpurun_out/t_conv.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/t_conv.log
{ echo "== weights resident"; timeout 300 python scripts/bench_gemm.py conv 5; echo "== off"; B2C_TAPGEMM_BRES=0 timeout 300 python scripts/bench_gemm.py conv 5; } > gpurun_out/conv.txt 2>&1; cat gpurun_out/conv.txt
