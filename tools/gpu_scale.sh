#!/bin/bash
# Multi-GPU bench lines on one box: bash tools/gpu_scale.sh N [workloads...]  (torchrun, one rank per GPU, NCCL)
mkdir -p gpurun_out
N=$1; shift
for w in "$@"; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --workload $w --steps ${STEPS:-10} --warmup 3 \
    > gpurun_out/bench_${w}_n$N.json 2> gpurun_out/bench_${w}_n$N.err; echo "bench $w n=$N rc=$?"
  tail -1 gpurun_out/bench_${w}_n$N.json | cut -c1-260
done
