#!/bin/bash
# Developer tool: validate a library variant (build/lib<name>.so) with the bit-exact GPU suites, then A/B it against the in-tree library.
#   bash tools/gpu_variant_check.sh <name> [more names for the A/B only]
mkdir -p gpurun_out build
cp tf_ode_b200/csrc/libtfode_b200.so build/libbase.so
N=$1
TFODE_LIB=$PWD/build/lib$N.so timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py tests/test_gpu_model_tree.py -m gpu -q -x -k "not long_horizon" > gpurun_out/pytest_$N.log 2>&1; echo "pytest($N) rc=$?"; tail -4 gpurun_out/pytest_$N.log
echo "== config3 early"; python tools/ab_bench.py --rounds 2 --workload config3 --pre 10 --steps 6 base "$@"
echo "== config3 late";  python tools/ab_bench.py --rounds 1 --workload config3 --pre 60 --steps 6 base "$@"
echo "== config2 groove"; python tools/ab_bench.py --rounds 1 --workload config2 --pre 42 --steps 6 base "$@"
echo "== config4"; python tools/ab_bench.py --rounds 1 --workload config4 base "$@"
echo "== config5"; python tools/ab_bench.py --rounds 1 --workload config5 --steps 3 base "$@"
