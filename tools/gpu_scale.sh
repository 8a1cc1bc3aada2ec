#!/bin/bash
# Multi-GPU bench lines on one box, launched as the driver launches them: bash tools/gpu_scale.sh N [N2 ...]   (default workload: the config-5 sweep)
mkdir -p gpurun_out
nvidia-smi -L | head -8
for N in "$@"; do
  timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps ${STEPS:-20} --warmup ${WARMUP:-5} ${EXTRA} \
    > gpurun_out/bench_config5_n$N.json 2> gpurun_out/bench_config5_n$N.err; echo "bench n=$N rc=$?"
  python -c "import json;d=json.loads(open('gpurun_out/bench_config5_n$N.json').read().strip().splitlines()[-1]);print('n=$N value %.2fM e2e %.2fM ms %.1f parity %s scaling %s'%(d['value']/1e6,d['e2e']['value']/1e6,d['ms_per_step'],d['parity']['bit_exact'],d['scaling']))" || tail -5 gpurun_out/bench_config5_n$N.err
done
