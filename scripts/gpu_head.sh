#!/bin/bash
# PatchGAN head kernels: parity, then the mma.sync kernels against the CUDA-core ones (B2C_HEAD_MMA=0)
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -k "test_head" > gpurun_out/t_head.log 2>&1; tail -3 gpurun_out/t_head.log
{ echo "== mma"; timeout 300 python scripts/bench_head.py; echo "== old"; B2C_HEAD_MMA=0 timeout 300 python scripts/bench_head.py; echo "== mma 64x32"; timeout 300 python scripts/bench_head.py 64 32; } > gpurun_out/head.txt 2>&1; cat gpurun_out/head.txt
