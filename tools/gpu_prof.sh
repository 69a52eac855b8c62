#!/bin/bash
# Development aid: ncu launch list + one full capture of the hot kernels on the N=1 bench workload (tools/prof_step.py).
mkdir -p gpurun_out
ARGS="$@"
timeout 300 python tools/prof_step.py $ARGS > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv python tools/prof_step.py $ARGS > gpurun_out/ncu_list.log 2>&1
timeout 300 python tools/prof_step.py $ARGS > gpurun_out/plain2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'field_kernel|tri_kernel|vertex_kernel|color_kernel' -s 4 -c 4 -o gpurun_out/prof -f python tools/prof_step.py $ARGS > gpurun_out/ncu_full.log 2>&1
tail -n 3 gpurun_out/ncu_list.log; tail -n 3 gpurun_out/ncu_full.log; cat gpurun_out/plain.log
