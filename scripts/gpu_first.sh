#!/bin/bash
# first-layer kernels: parity, then timings (packer groups / ring depths)
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -k "first_layer" > gpurun_out/t_first.log 2>&1; tail -3 gpurun_out/t_first.log
B2C_FIRST_NG=2 timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -k "first_layer" > gpurun_out/t_first2.log 2>&1; tail -2 gpurun_out/t_first2.log
{
echo "== default"; timeout 300 python scripts/bench_first.py 64 256 10
echo "== NG=2"; B2C_FIRST_NG=2 timeout 300 python scripts/bench_first.py 64 256 10
echo "== NG=3 NS=6"; B2C_FIRST_NS=6 timeout 300 python scripts/bench_first.py 64 256 10
echo "== 512"; timeout 300 python scripts/bench_first.py 16 512 10
} > gpurun_out/first.txt 2>&1
grep -v "first_dgrad" gpurun_out/first.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"first_.*_tma" -c 6 -f -o gpurun_out/ncu_first_tma3 python scripts/bench_first.py 64 256 1 > gpurun_out/ncu_first_tma3.log 2>&1
ls -la gpurun_out/ncu_first_tma3.ncu-rep
