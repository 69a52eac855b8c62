#!/bin/bash
# Development aid: one GPU-box session = gpu tests, smoke, bench, ncu launch list + full captures (N = 1 workload in both semantics and
# rank 0's slab of the N = 8 workload). Outputs under gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 6 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -n 2 gpurun_out/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
tail -c 1500 gpurun_out/bench.log; tail -n 3 gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2> gpurun_out/bench_ref.err; tail -c 600 gpurun_out/bench_ref.log
if [ "$1" = "ncu" ]; then
  K='field_kernel|tri_kernel|vertex_kernel|color_kernel'
  timeout 300 python tools/prof_step.py > gpurun_out/plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches.csv python tools/prof_step.py > gpurun_out/ncu_list.log 2>&1
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"$K" -s 4 -c 4 -o gpurun_out/prof -f python tools/prof_step.py > gpurun_out/ncu_full.log 2>&1
  timeout 300 python tools/prof_step.py --native > gpurun_out/plain_native.log 2>&1 &&
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"$K" -s 4 -c 4 -o gpurun_out/prof_native -f python tools/prof_step.py --native > gpurun_out/ncu_full_native.log 2>&1
  timeout 300 python tools/prof_step.py --slab=0/8 > gpurun_out/plain_slab.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_slab0of8.csv python tools/prof_step.py --slab=0/8 > gpurun_out/ncu_list_slab.log 2>&1
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"$K" -s 4 -c 4 -o gpurun_out/prof_slab -f python tools/prof_step.py --slab=0/8 > gpurun_out/ncu_full_slab.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'prune_fused' -s 2 -c 2 -o gpurun_out/prof_cull -f python tools/prof_step.py --slab=0/8 > gpurun_out/ncu_full_cull.log 2>&1
  tail -n 2 gpurun_out/ncu_list.log gpurun_out/ncu_full.log gpurun_out/ncu_full_native.log gpurun_out/ncu_full_slab.log; cat gpurun_out/plain.log | cut -c1-300
fi
