#!/bin/bash
# tcgen05 actor kernel: its parity tests, then the stand-alone timing (tools/policy_bench.py) and phase timelines -- each under its own timeout.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -m gpu -q -k "policy_tc or tc_actor" > gpurun_out/pytest_policy_tc.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_policy_tc.log
timeout 120 python tools/policy_bench.py 16384 > gpurun_out/policy_bench.log 2>&1; echo "bench rc=$?"; cat gpurun_out/policy_bench.log | cut -c1-250
timeout 120 python tools/policy_bench.py 131072 > gpurun_out/policy_bench_131072.log 2>&1; cut -c1-120 gpurun_out/policy_bench_131072.log
bash tools/gpu_policy_tl.sh
