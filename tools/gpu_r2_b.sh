#!/bin/bash
# Round-2 GPU visit B (2 GPUs): NCCL rollout test, 2-GPU bench line (config 5 strong scaling + K7 gather), then single-GPU profiling of config 3:
# phase split, ncu launch list, ncu --set full of step_kernel.
mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m pytest tests/test_gpu_nccl_rollout.py tests/test_td3.py -m gpu -q -x > gpurun_out/pytest_nccl.log 2>&1; echo "pytest nccl rc=$?"; tail -5 gpurun_out/pytest_nccl.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 \
  > gpurun_out/bench_config5_n2.json 2> gpurun_out/bench_config5_n2.err; echo "bench n2 rc=$?"; cat gpurun_out/bench_config5_n2.json; tail -3 gpurun_out/bench_config5_n2.err
export CUDA_VISIBLE_DEVICES=0
python tools/phase_timing.py config3 2>&1 | grep -v "warning\|pragma\|\^\|Remark\|^$" > gpurun_out/phase_config3.log; cat gpurun_out/phase_config3.log
B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-verify --no-actor-tc"
timeout 300 $B > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_config3.csv $B > gpurun_out/ncu1.log 2>&1
echo "ncu list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 5 -c 1 -o gpurun_out/prof_step_config3 -f $B > gpurun_out/ncu2.log 2>&1
echo "ncu full step rc=$?"; tail -2 gpurun_out/ncu2.log | cut -c1-200
