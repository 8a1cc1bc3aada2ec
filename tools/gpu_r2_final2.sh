#!/bin/bash
# Round-2 final numbers of the committed library: bench lines (config 3 default with CPU arms, reference arm, config 2, config 4), launch list + full ncu capture.
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench_config3.json 2> gpurun_out/bench_config3.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench_config3.json; tail -3 gpurun_out/bench_config3.err
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref_config3.json 2>&1; cut -c1-300 gpurun_out/bench_ref_config3.json
for w in config2 config4; do
  timeout 900 python bench.py --workload $w > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w rc=$?"
  python -c "import json;d=json.load(open('gpurun_out/bench_$w.json'));print('$w value %.3fM e2e %.3fM ms %.2f cpu %.3fM parity %s'%(d['value']/1e6,d['e2e']['value']/1e6,d['ms_per_step'],d['cpu_baseline']['value']/1e6,d['parity']['bit_exact']))"
done
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-verify --no-actor-tc"
timeout 300 $B > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_config3.csv $B > gpurun_out/ncu1.log 2>&1
echo "ncu list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 5 -c 1 -o gpurun_out/prof_step_config3 -f $B > gpurun_out/ncu2.log 2>&1
echo "ncu full step rc=$?"; tail -2 gpurun_out/ncu2.log | cut -c1-200
timeout 300 python tools/policy_bench.py > gpurun_out/policy_bench.log 2>&1; cat gpurun_out/policy_bench.log
