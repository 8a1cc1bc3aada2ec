#!/bin/bash
# ncu capture of policy_tc_kernel (one launch, --set full) after the plain run has exited 0
mkdir -p gpurun_out
timeout 120 python tools/policy_bench.py 16384 > gpurun_out/policy_bench.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/policy_bench.log; exit 1; }
cut -c1-200 gpurun_out/policy_bench.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:policy_tc_kernel -s 3 -c 1 -o gpurun_out/prof_policy_tc -f python tools/policy_bench.py 16384 > gpurun_out/ncu_policy.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_policy.log | cut -c1-200
timeout 60 tools/policy_tc_timeline_base 16384 > gpurun_out/tl_base_16384.txt 2>&1; head -1 gpurun_out/tl_base_16384.txt
