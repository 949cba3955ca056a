#!/bin/bash
# first-layer kernels: parity, then timings
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -k "first_layer" > gpurun_out/t_first.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/t_first.log
{ echo "== default"; timeout 300 python scripts/bench_first.py 64 256 10; echo "== old dgrad"; B2C_FIRST_DGRAD2=0 timeout 300 python scripts/bench_first.py 64 256 10 | grep dgrad; echo "== 512"; timeout 300 python scripts/bench_first.py 16 512 10 | grep dgrad; } > gpurun_out/first.txt 2>&1
cat gpurun_out/first.txt
