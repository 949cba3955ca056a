#!/bin/bash
# Evidence for profiles/: launch list of one step, ncu --set full of its GEMM and bandwidth kernels, kprof, bench lines.
mkdir -p gpurun_out/prof
export PYTHONUNBUFFERED=1
P=gpurun_out/prof
timeout 600 python scripts/kprof.py 64 256 3 > $P/kprof_step.txt 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file $P/launches.csv python scripts/one_step.py > $P/launches.log 2>&1
timeout 1500 ncu --set full --clock-control none --profile-from-start off -k regex:"tapgemm|convt64|wgrad_kernel" -f -o $P/gemm_step python scripts/one_step.py > $P/gemm_step.log 2>&1
timeout 1500 ncu --set full --clock-control none --profile-from-start off -k regex:"plane_|first_|head_|multi_adam|lrelu|strided|sn_" -f -o $P/bw_step python scripts/one_step.py > $P/bw_step.log 2>&1
ncu -i $P/gemm_step.ncu-rep --page raw --csv > $P/gemm_step_raw.csv 2>/dev/null
ncu -i $P/bw_step.ncu-rep --page raw --csv > $P/bw_step_raw.csv 2>/dev/null
rm -f $P/gemm_step.ncu-rep $P/bw_step.ncu-rep
ls -la $P
