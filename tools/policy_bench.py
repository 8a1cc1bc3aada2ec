"""Developer tool: times the actor kernels alone (tfode_policy_infer on m rows), tcgen05 vs fp32 CUDA-core path.

  python tools/policy_bench.py [m]      (on the GPU box)
Prints ms per call, TFLOP/s (543 744 flop per row) and the max action difference between the paths.
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tf_ode_b200 import RoverBatch, abi, glorot_policy, make_config  # noqa: E402
import bench  # noqa: E402  (ClockSampler: nvidia-smi clocks and throttle reasons during the timed loops)

m = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
x = torch.from_numpy((np.random.RandomState(3).rand(m, 17).astype(np.float32) * 2 - 1)).cuda()
ys = {}
for name, prec in (("tcgen05", abi.POLICY_TC), ("fp32", abi.POLICY_FP32)):
    g = RoverBatch(make_config(abi.PRESET_TRAINING, 8, policy_precision=prec), stream=torch.cuda.current_stream().cuda_stream)
    g.load_policy(glorot_policy())
    y = torch.zeros(m, 2, device="cuda")
    for _ in range(5):
        g.policy_infer(x, y)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 20000 if prec == abi.POLICY_TC else 2000   # long enough (about 0.5 s) for nvidia-smi to sample the clocks under this load
    clk = bench.ClockSampler(0); clk.mark()
    a.record()
    for _ in range(reps):
        g.policy_infer(x, y)
    b.record()
    torch.cuda.synchronize()
    clocks = clk.stop()
    ms = a.elapsed_time(b) / reps
    ys[name] = y.cpu().numpy()
    print("%-8s m=%d  %.4f ms/call  %.2f TFLOP/s  clocks %s" % (name, m, ms, m * 543744 / (ms * 1e-3) / 1e12, clocks))
    g.close()
print("max |tcgen05 - fp32| = %.3e" % np.abs(ys["tcgen05"] - ys["fp32"]).max())
