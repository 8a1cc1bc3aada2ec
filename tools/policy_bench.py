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
side = torch.cuda.Stream()
for name, prec in (("tcgen05", abi.POLICY_TC), ("fp32", abi.POLICY_FP32)):
    # One ctypes call costs the host ~10 us, as much as the tcgen05 kernel runs: the launches are captured into a CUDA graph (tfode_policy_infer
    # only enqueues a kernel on the handle's stream) and the graph is replayed, so the figure is the GPU's back-to-back rate, not Python's.
    with torch.cuda.stream(side):
        g = RoverBatch(make_config(abi.PRESET_TRAINING, 8, policy_precision=prec), stream=side.cuda_stream)
        g.load_policy(glorot_policy())
        y = torch.zeros(m, 2, device="cuda")
        for _ in range(5):
            g.policy_infer(x, y)
        side.synchronize()
        per_graph = 100 if prec == abi.POLICY_TC else 10
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            for _ in range(per_graph):
                g.policy_infer(x, y)
        graph.replay(); side.synchronize()
        replays = max(1, int(0.6 / (per_graph * (20e-6 if prec == abi.POLICY_TC else 400e-6) * m / 16384)))   # about 0.6 s: nvidia-smi samples the clocks under this load
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        clk = bench.ClockSampler(0); clk.mark()
        a.record(side)
        for _ in range(replays):
            graph.replay()
        b.record(side)
        side.synchronize()
        clocks = clk.stop()
        ms = a.elapsed_time(b) / (replays * per_graph)
        ys[name] = y.cpu().numpy()
        print("%-8s m=%d  %.4f ms/call  %.2f TFLOP/s  (%d graph replays x %d launches)  clocks %s" % (name, m, ms, m * 543744 / (ms * 1e-3) / 1e12, replays, per_graph, clocks))
        del graph
        g.close()
print("max |tcgen05 - fp32| = %.3e" % np.abs(ys["tcgen05"] - ys["fp32"]).max())
