#!/bin/bash
# Round-2 GPU visit D: the per-scene lock-step unroll (in-tree library) through the bit-exact suites, the helper-loop variant on top, A/B of both.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x -k "not long_horizon" > gpurun_out/pytest_gpu_d.log 2>&1; echo "pytest(in-tree) rc=$?"; tail -3 gpurun_out/pytest_gpu_d.log
TFODE_LIB=$PWD/build/libhloop.so timeout 1200 python -m pytest tests -m gpu -q -x -k "not long_horizon" > gpurun_out/pytest_hloop.log 2>&1; echo "pytest(hloop) rc=$?"; tail -3 gpurun_out/pytest_hloop.log
bash tools/gpu_ab_all.sh hloop
