#!/bin/bash
# first-layer kernels: parity of both variants, then timings (register gather vs TMA-staged, ring depths)
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -k "first_layer" > gpurun_out/t_first.log 2>&1; tail -5 gpurun_out/t_first.log
B2C_FIRST_TMA=0 timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -k "first_layer" > gpurun_out/t_first_old.log 2>&1; tail -2 gpurun_out/t_first_old.log
{
echo "== register gather"; B2C_FIRST_TMA=0 timeout 300 python scripts/bench_first.py 64 256 10
echo "== TMA default NS"; timeout 300 python scripts/bench_first.py 64 256 10
for ns in 2 3 4 6; do echo "== TMA NS=$ns"; B2C_FIRST_NS=$ns timeout 300 python scripts/bench_first.py 64 256 10; done
echo "== TMA 512"; timeout 300 python scripts/bench_first.py 16 512 10
} > gpurun_out/first.txt 2>&1
cat gpurun_out/first.txt
