#!/bin/bash
mkdir -p gpurun_out
{
for shape in "64 256 512" "64 1024 256" "64 4096 128" "16 16384 128" "16 4096 256"; do
  for cs in 0 1 2 4 8; do
    echo "== shape $shape CS=$cs"; B2C_NORM_CS=$cs timeout 120 python scripts/bench_norm.py $shape 10 2>&1 | grep -E "in_act_bwd|in_act_fwd"
  done
done
} > gpurun_out/norm_cs.txt 2>&1
cat gpurun_out/norm_cs.txt
