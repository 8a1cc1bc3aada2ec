#!/bin/bash
# ncu --set full capture of one step_kernel launch of the default workload (config 3), after the plain command exited 0
mkdir -p gpurun_out
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-verify --no-actor-tc"
timeout 300 $B > gpurun_out/plain.log 2>&1 || { echo plain failed; exit 1; }
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 5 -c 1 -o gpurun_out/prof_step_config3 -f $B > gpurun_out/ncu2.log 2>&1
echo "ncu full step rc=$?"; tail -2 gpurun_out/ncu2.log | cut -c1-200
