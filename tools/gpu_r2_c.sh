#!/bin/bash
# Round-2 GPU visit C: Dantzig path, TD3 loop and bit-exact suites with the step-box culling; A/B of the culling; Dantzig bench line.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 -k "not long_horizon" > gpurun_out/pytest_gpu_c.log 2>&1; echo "pytest rc=$?"; tail -14 gpurun_out/pytest_gpu_c.log
echo "== config3 early"; python tools/ab_bench.py --rounds 2 --workload config3 --pre 10 nocull cull
echo "== config3 late";  python tools/ab_bench.py --rounds 1 --workload config3 --pre 60 --steps 6 nocull cull
echo "== config2"; python tools/ab_bench.py --rounds 1 nocull cull
python tools/phase_timing.py config3 2>&1 | grep -v "warning\|pragma\|\^\|Remark\|^$" > gpurun_out/phase_config3_cull.log; cat gpurun_out/phase_config3_cull.log
timeout 900 python bench.py --solver dantzig --envs 4096 --steps 4 --no-cpu-baseline --no-actor-tc > gpurun_out/bench_config3_dantzig.json 2> gpurun_out/bench_config3_dantzig.err; echo "bench dantzig rc=$?"; cut -c1-1500 gpurun_out/bench_config3_dantzig.json; tail -3 gpurun_out/bench_config3_dantzig.err
