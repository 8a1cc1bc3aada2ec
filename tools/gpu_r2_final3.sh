#!/bin/bash
# Final round-2 validation of the committed library: whole GPU suite, smoke, default bench + reference arm, tcgen05 actor timing / ncu / timeline, launch list.
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q --durations=6 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -12 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_config3.json 2> gpurun_out/bench_config3.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/bench_config3.json; tail -3 gpurun_out/bench_config3.err
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref_config3.json 2>&1; cut -c1-200 gpurun_out/bench_ref_config3.json
timeout 120 python tools/policy_bench.py 16384 > gpurun_out/policy_bench.log 2>&1; echo "policy bench rc=$?"; cut -c1-220 gpurun_out/policy_bench.log
timeout 120 python tools/policy_bench.py 131072 > gpurun_out/policy_bench_131072.log 2>&1; cut -c1-120 gpurun_out/policy_bench_131072.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:policy_tc_kernel -s 3 -c 1 -o gpurun_out/prof_policy_tc -f python tools/policy_bench.py 16384 > gpurun_out/ncu_policy.log 2>&1; echo "ncu policy rc=$?"
timeout 60 tools/policy_tc_timeline_base 16384 > gpurun_out/tl_base_16384.txt 2>&1; head -1 gpurun_out/tl_base_16384.txt
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-verify --no-actor-tc"
timeout 300 $B > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_config3.csv $B > gpurun_out/ncu1.log 2>&1
echo "ncu list rc=$?"
