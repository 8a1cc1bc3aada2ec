#!/bin/bash
# Round-2 GPU visit E: the persistent (group, slice) launch -- equality with the plain launch first (short timeout: a scheduling bug would hang), then
# the bit-exact suites, then timing with and without it.
mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_gpu_configs.py -m gpu -q -x -k "persistent" > gpurun_out/pytest_persist.log 2>&1; rc=$?; echo "pytest(persistent) rc=$rc"; tail -5 gpurun_out/pytest_persist.log
if [ $rc -ne 0 ]; then exit 1; fi
timeout 1200 python -m pytest tests -m gpu -q -x -k "not long_horizon" > gpurun_out/pytest_gpu_e.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_e.log
cp tf_ode_b200/csrc/libtfode_b200.so build/libpersist.so
for w in config2 config4 config3; do
  for pre in 3 42; do
    echo "== $w pre=$pre"; TFODE_PERSISTENT=0 python tools/ab_bench.py --rounds 1 --workload $w --pre $pre --steps 6 persist | tail -1; TFODE_PERSISTENT=1 python tools/ab_bench.py --rounds 1 --workload $w --pre $pre --steps 6 persist | tail -1
  done
done
for w in config2 config4; do
  timeout 600 python bench.py --workload $w --no-cpu-baseline > gpurun_out/bench_${w}_persist.json 2> gpurun_out/bench_${w}_persist.err; echo "bench $w rc=$?"
  python -c "import json;d=json.load(open('gpurun_out/bench_${w}_persist.json'));print('$w value %.3fM e2e %.3fM ms %.2f parity %s'%(d['value']/1e6,d['e2e']['value']/1e6,d['ms_per_step'],d['parity']['bit_exact']))"
done
