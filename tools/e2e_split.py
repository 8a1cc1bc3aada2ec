"""Where the e2e arm's time goes (config 2): per-step device time of tfode_batch_step alone over a long run (the workload drifts as the
rovers reach the groove), then per-step wall and device time of tfode_batch_rollout_host, interleaved with plain steps at the same state."""
import os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from tf_ode_b200 import RoverBatch

n = 4096
cfg = bench.make_cfg("config2", n, env_id_offset=0, device=0)
g = RoverBatch(cfg, stream=torch.cuda.current_stream().cuda_stream)
S = 100
def dev_ms(fn):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); fn(); b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b)
steps = [dev_ms(lambda: g.step(S)) for _ in range(60)]
print("step-only ms, steps 0..59:", " ".join("%.2f" % x for x in steps))
act = torch.zeros(n, 2).pin_memory().numpy(); obs = torch.zeros(n, 17).pin_memory().numpy()
rew = torch.zeros(n).pin_memory().numpy(); done = torch.zeros(n, dtype=torch.uint8).pin_memory().numpy()
for k in range(10):
    s = dev_ms(lambda: g.step(S))
    t0 = time.perf_counter(); d = dev_ms(lambda: g.rollout_host(act, S, obs, rew, done)); w = (time.perf_counter() - t0) * 1e3
    t0 = time.perf_counter(); g.rollout_host(act, S, obs, rew, done); w2 = (time.perf_counter() - t0) * 1e3
    print("pair %d: step-only %.2f ms | rollout_host device %.2f ms wall(with events) %.2f ms | rollout_host wall %.2f ms" % (k, s, d, w, w2))
