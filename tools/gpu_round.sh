#!/bin/bash
# Development aid: one GPU-box session = gpu tests, bench, ncu launch list + one full capture. Outputs under gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -n 2 gpurun_out/smoke.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
tail -c 2500 gpurun_out/bench.log; tail -5 gpurun_out/bench.err
if [ "$1" = "ncu" ]; then
  timeout 300 python tools/prof_step.py > gpurun_out/plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv python tools/prof_step.py > gpurun_out/ncu_list.log 2>&1
  timeout 300 python tools/prof_step.py > gpurun_out/plain2.log 2>&1 &&
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'field_kernel|tri_kernel|vertex_kernel|color_kernel' -s 4 -c 4 -o gpurun_out/prof -f python tools/prof_step.py > gpurun_out/ncu_full.log 2>&1
  tail -3 gpurun_out/ncu_list.log gpurun_out/ncu_full.log; cat gpurun_out/plain.log
fi
if [ "$1" = "ncu" ]; then
  timeout 300 python tools/prof_step.py --native > gpurun_out/plain_native.log 2>&1 &&
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'field_kernel|vertex_kernel' -s 2 -c 2 -o gpurun_out/prof_native -f python tools/prof_step.py --native > gpurun_out/ncu_full_native.log 2>&1
  tail -2 gpurun_out/ncu_full_native.log
fi
