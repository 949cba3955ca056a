#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
B2C_TAPGEMM_2CTA=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"tapgemm_p" -c 1 -f -o gpurun_out/ncu_pair2 python scripts/bench_gemm.py conv 1 > gpurun_out/ncu_pair2.log 2>&1
ls -la gpurun_out/ncu_pair2.ncu-rep
