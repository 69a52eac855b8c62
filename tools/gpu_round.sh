#!/bin/bash
# Development aid: one GPU-box session = quick shake-out, gpu tests, bench, launch list. Outputs under gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.txt 2>&1
timeout 300 python tools/gpu_quick.py > gpurun_out/quick.log 2>&1; echo "quick rc=$?" >> gpurun_out/quick.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
tail -5 gpurun_out/quick.log; tail -3 gpurun_out/pytest_gpu.log; tail -c 3000 gpurun_out/bench.log; tail -5 gpurun_out/bench.err
