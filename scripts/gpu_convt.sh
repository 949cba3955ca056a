#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
{ echo "== 8 epilogue warps"; timeout 300 python scripts/bench_gemm.py convt 5
  echo "== 4 epilogue warps (previous build)"; cp gan-calibrated-semi-supervised-learning_b200/b200cgan/libb2c.so /tmp/new.so; cp scripts/probe/libb2c_old4w.so gan-calibrated-semi-supervised-learning_b200/b200cgan/libb2c.so; timeout 300 python scripts/bench_gemm.py convt 5; cp /tmp/new.so gan-calibrated-semi-supervised-learning_b200/b200cgan/libb2c.so; } > gpurun_out/convt.txt 2>&1; cat gpurun_out/convt.txt
