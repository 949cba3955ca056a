#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -k "test_head" > gpurun_out/t_head.log 2>&1; tail -3 gpurun_out/t_head.log
{ echo "== new"; timeout 300 python scripts/bench_head.py; echo "== new 64x32"; timeout 300 python scripts/bench_head.py 64 32; } > gpurun_out/head.txt 2>&1; cat gpurun_out/head.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"head_.*2_kernel" -c 3 -f -o gpurun_out/ncu_head2 python scripts/bench_head.py 64 16 0 > gpurun_out/ncu_head2.log 2>&1
ls -la gpurun_out/ncu_head2.ncu-rep
