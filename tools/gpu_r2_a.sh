#!/bin/bash
# Round-2 GPU visit A: full GPU test suite, smoke, A/B of the row-layout / pairing variants against the round-1 library, bench lines.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=10 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
V="$@"
echo "== config2 cheap segment"; python tools/ab_bench.py --rounds 2 $V
echo "== config2 groove segment"; python tools/ab_bench.py --rounds 2 --pre 42 --steps 6 $V
echo "== config3"; python tools/ab_bench.py --rounds 1 --workload config3 --pre 10 $V
echo "== config4"; python tools/ab_bench.py --rounds 1 --workload config4 $V
echo "== config5"; python tools/ab_bench.py --rounds 1 --workload config5 --steps 3 $V
timeout 900 python bench.py > gpurun_out/bench_config3.json 2> gpurun_out/bench_config3.err; echo "bench rc=$?"; cat gpurun_out/bench_config3.json; tail -3 gpurun_out/bench_config3.err
timeout 900 python bench.py --workload config2 > gpurun_out/bench_config2.json 2> gpurun_out/bench_config2.err; echo "bench config2 rc=$?"; cut -c1-400 gpurun_out/bench_config2.json; tail -3 gpurun_out/bench_config2.err
python tools/phase_timing.py config2 2>&1 | grep -v "warning\|pragma\|\^\|Remark\|^$" > gpurun_out/phase_config2.log; cat gpurun_out/phase_config2.log
