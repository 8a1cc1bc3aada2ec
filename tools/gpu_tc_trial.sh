#!/bin/bash
# First visit for the tcgen05 actor: its tests under a short timeout (a wrong descriptor can hang the kernel), then timing.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -q -x -k "policy" > gpurun_out/pytest_tc.log 2>&1; echo "pytest policy rc=$?"; tail -15 gpurun_out/pytest_tc.log
timeout 120 python tools/policy_bench.py > gpurun_out/policy_bench.log 2>&1; echo "policy_bench rc=$?"; cat gpurun_out/policy_bench.log | tail -5
