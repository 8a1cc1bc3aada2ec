#!/bin/bash
# One GPU-box visit: tests, smoke, bench lines of every workload, CPU arm, and (with "ncu") the launch list plus one
# full capture each of step_kernel (config 2) and policy_tc_kernel (16384 rows).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_config2.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench_config2.json; tail -3 gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 5 > gpurun_out/bench_ref_config2.json 2>&1; cat gpurun_out/bench_ref_config2.json
for w in config3 config4 config5; do
  timeout 600 python bench.py --workload $w --steps 10 --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w rc=$?"; cut -c1-200 gpurun_out/bench_$w.json
done
timeout 120 python tools/policy_bench.py > gpurun_out/policy_bench.log 2>&1; cat gpurun_out/policy_bench.log
if [ "$1" = "ncu" ]; then
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_config2.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 4 -c 1 -o gpurun_out/prof_step_config2 -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
echo "ncu full step rc=$?"; tail -2 gpurun_out/ncu2.log | cut -c1-200
timeout 300 python bench.py --workload config3 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain3.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_config3.csv python bench.py --workload config3 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu3.log 2>&1
echo "ncu list config3 rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:policy_tc_kernel -s 5 -c 1 -o gpurun_out/prof_policy_tc -f python tools/policy_bench.py > gpurun_out/ncu4.log 2>&1
echo "ncu full policy rc=$?"; tail -2 gpurun_out/ncu4.log
fi
