#!/bin/bash
# Round-2 final single-GPU visit: whole GPU suite, smoke, bench lines of every workload, CPU arm, launch list + full capture of step_kernel (config 3).
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x --durations=10 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -16 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_config3.json 2> gpurun_out/bench_config3.err; echo "bench rc=$?"; cat gpurun_out/bench_config3.json; tail -3 gpurun_out/bench_config3.err
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref_config3.json 2>&1; cut -c1-600 gpurun_out/bench_ref_config3.json
for w in config2 config4; do
  timeout 900 python bench.py --workload $w > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w rc=$?"
  python -c "import json;d=json.load(open('gpurun_out/bench_$w.json'));print('$w value %.3fM e2e %.3fM ms %.2f cpu %.3fM parity %s seg %s'%(d['value']/1e6,d['e2e']['value']/1e6,d['ms_per_step'],d['cpu_baseline']['value']/1e6,d['parity']['bit_exact'],d['segments']))"
done
timeout 1500 python bench.py --workload config5 --steps 10 --no-cpu-baseline > gpurun_out/bench_config5_n1.json 2> gpurun_out/bench_config5_n1.err; echo "bench config5 n1 rc=$?"
python -c "import json;d=json.load(open('gpurun_out/bench_config5_n1.json'));print('config5 n1 value %.3fM e2e %.3fM ms %.2f parity %s'%(d['value']/1e6,d['e2e']['value']/1e6,d['ms_per_step'],d['parity']))"
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-verify --no-actor-tc"
timeout 300 $B > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_config3.csv $B > gpurun_out/ncu1.log 2>&1
echo "ncu list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 5 -c 1 -o gpurun_out/prof_step_config3 -f $B > gpurun_out/ncu2.log 2>&1
echo "ncu full step rc=$?"; tail -2 gpurun_out/ncu2.log | cut -c1-200
