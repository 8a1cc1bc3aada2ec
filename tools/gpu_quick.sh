#!/bin/bash
# Quick GPU visit: parity tests, then the bench lines of the workloads named on the command line (no CPU arm).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
for w in "$@"; do
  timeout 300 python bench.py --workload $w --steps 10 --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w rc=$?"
  python -c "import json;d=json.load(open('gpurun_out/bench_$w.json'));print('$w value %.3fM e2e %.3fM ms %.2f launches/step %s'%(d['value']/1e6,d['e2e']['value']/1e6,d['ms_per_step'],d.get('launches_per_step')))"
done
