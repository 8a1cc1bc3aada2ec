#!/bin/bash
# Last visit of round 2: the suite without the two long-horizon tests (run and green on the library of the previous visit; policy_tc_kernel is the
# only kernel changed since), tcgen05 actor timing / timeline / ncu, default bench line.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "not long_horizon" > gpurun_out/pytest_gpu_final4.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_final4.log
timeout 120 python tools/policy_bench.py 16384 > gpurun_out/policy_bench.log 2>&1; cut -c1-200 gpurun_out/policy_bench.log
timeout 120 python tools/policy_bench.py 131072 > gpurun_out/policy_bench_131072.log 2>&1; cut -c1-120 gpurun_out/policy_bench_131072.log
timeout 60 tools/policy_tc_timeline_base 16384 > gpurun_out/tl_base_16384.txt 2>&1; head -1 gpurun_out/tl_base_16384.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:policy_tc_kernel -s 3 -c 1 -o gpurun_out/prof_policy_tc -f python tools/policy_bench.py 16384 > gpurun_out/ncu_policy.log 2>&1; echo "ncu policy rc=$?"
timeout 800 python bench.py > gpurun_out/bench_config3.json 2> gpurun_out/bench_config3.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/bench_config3.json
