#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"plane_cluster_kernel" -c 2 -f -o gpurun_out/ncu_cluster python scripts/bench_norm.py 64 4096 128 1 > gpurun_out/ncu_cluster.log 2>&1
ls -la gpurun_out/ncu_cluster.ncu-rep
