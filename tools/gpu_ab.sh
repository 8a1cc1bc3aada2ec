#!/bin/bash
# Developer tool: parity subset + same-box A/B of two library variants (tools/build_variant.sh) + phase split of the working tree.
#   tools/gpu_ab.sh <old> <new>
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_model_tree.py -m gpu -q -x -k "bit_exact or contact_counts or config2 or terrain" > gpurun_out/pytest_ab.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_ab.log
python tools/ab_bench.py --rounds 2 $1 $2
python tools/ab_bench.py --rounds 1 --workload config3 $1 $2
python tools/ab_bench.py --rounds 1 --workload config4 $1 $2
python tools/ab_bench.py --rounds 1 --workload config5 --steps 3 $1 $2
python tools/phase_timing.py config2 2>&1 | grep -v "warning\|pragma\|\^\|Remark\|^$" > gpurun_out/phase_ab.log; cat gpurun_out/phase_ab.log
