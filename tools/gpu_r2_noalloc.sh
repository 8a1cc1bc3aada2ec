#!/bin/bash
# The no-alloc entry point's test alone in a fresh process, then the whole GPU suite (which collects it last), smoke, and the
# committed library's default bench lines.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -4 gpurun_out/pytest_noalloc_alone.log
timeout 2400 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -14 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_config3.json 2> gpurun_out/bench_config3.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench_config3.json; tail -3 gpurun_out/bench_config3.err
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref_config3.json 2>&1; cut -c1-300 gpurun_out/bench_ref_config3.json
