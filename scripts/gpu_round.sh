#!/bin/bash
# One GPU session: kernel tests first (a hang there must not cost the model tests), then model tests, kprof, bench.
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 900 python -m pytest tests/test_abi.py tests/test_kernels_gpu.py tests/test_kernels2_gpu.py -m gpu -q --timeout 300 -x > gpurun_out/t_kernels.log 2>&1
echo "kernels rc=$?" >> gpurun_out/summary.txt
tail -5 gpurun_out/t_kernels.log
timeout 1500 python -m pytest tests/test_dlayers_gpu.py tests/test_layers_gpu.py tests/test_models_gpu.py tests/test_graph_gpu.py tests/test_sizes_gpu.py -m gpu -q --timeout 600 -s > gpurun_out/t_models.log 2>&1
echo "models rc=$?" >> gpurun_out/summary.txt
tail -15 gpurun_out/t_models.log
timeout 600 python scripts/kprof.py 64 256 3 > gpurun_out/kprof.txt 2>&1
echo "kprof rc=$?" >> gpurun_out/summary.txt
head -40 gpurun_out/kprof.txt
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench rc=$?" >> gpurun_out/summary.txt
cat gpurun_out/bench.json | head -c 3000
cat gpurun_out/summary.txt
