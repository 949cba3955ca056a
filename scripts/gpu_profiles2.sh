#!/bin/bash
# Final-state evidence for profiles/: kprof, launch list of one step, one ncu pass of the whole step with the summary metrics
# (tensor-pipe active, DRAM bytes, L2 / DRAM throughput, registers), --set full captures of the two top kernels.
mkdir -p gpurun_out/prof
export PYTHONUNBUFFERED=1
P=gpurun_out/prof
timeout 600 python scripts/kprof.py 64 256 3 > $P/kprof_step.txt 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file $P/launches.csv python scripts/one_step.py > $P/launches.log 2>&1
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,lts__throughput.avg.pct_of_peak_sustained_elapsed,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,launch__registers_per_thread,launch__grid_size
timeout 1500 ncu --metrics $M --clock-control none --profile-from-start off --csv --page raw --log-file $P/step_raw.csv python scripts/one_step.py > $P/step.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"tapgemm_pair_kernel" -c 1 -f -o $P/top_conv python scripts/one_step.py > $P/top_conv.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"wgrad_pair_kernel" -c 1 -f -o $P/top_wgrad python scripts/one_step.py > $P/top_wgrad.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"convt64_kernel" -c 1 -f -o $P/top_convt64 python scripts/one_step.py > $P/top_convt64.log 2>&1
ls -la $P
