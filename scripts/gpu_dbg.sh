#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"first_.*_tma" -c 6 -f -o gpurun_out/ncu_first_tma python scripts/bench_first.py 64 256 1 > gpurun_out/ncu_first_tma.log 2>&1
ls -la gpurun_out/ncu_first_tma.ncu-rep
