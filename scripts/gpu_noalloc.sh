#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/t_noalloc.log 2>&1; echo "alone rc=$?"; tail -4 gpurun_out/t_noalloc.log
timeout 2400 python -m pytest tests -m gpu -x -q --timeout 900 > gpurun_out/t_all.log 2>&1; echo "all rc=$?"; tail -4 gpurun_out/t_all.log
