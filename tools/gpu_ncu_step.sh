#!/bin/bash
# One full ncu capture of step_kernel (config given as $1, default config2) with source correlation.
mkdir -p gpurun_out
W=${1:-config2}
timeout 300 python bench.py --workload $W --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 4 -c 1 -o gpurun_out/prof_$W -f python bench.py --workload $W --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
echo "ncu full rc=$?"; tail -3 gpurun_out/ncu2.log; ls -la gpurun_out/*.ncu-rep
