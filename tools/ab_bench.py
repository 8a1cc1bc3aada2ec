"""Developer tool: same-box A/B timing of library variants built by tools/build_variant.sh.

  python tools/ab_bench.py [--workload config2] [--rounds 3] [--steps 8] name1 name2 ...
Each measurement runs in its own process (fresh context); variants are interleaved round-robin so that box-to-box and
drift effects cancel.  Prints ms per bench step (100 sub-steps of every environment) per variant: all rounds and the best.
"""
import argparse
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

ONE = r'''
import sys, os
sys.path.insert(0, %r)
import torch
import tf_ode_b200.engine as eng
eng.LIB_PATH = sys.argv[1]
import bench
from tf_ode_b200 import RoverBatch, abi, glorot_policy
wl, steps, pre = sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
n = bench.dev_envs(wl)
cfg = bench.make_cfg(wl, n)
if os.path.basename(sys.argv[1]).startswith("libr1"):
    cfg.abi_version = 1   # the round-1 library: same config struct, older entry points
g = RoverBatch(cfg, stream=torch.cuda.current_stream().cuda_stream)
if g.cfg.controller == abi.CTRL_MLP:
    g.load_policy(glorot_policy())
for _ in range(pre):
    g.step(100)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(steps):
    g.step(100)
b.record()
torch.cuda.synchronize()
print("%%.3f" %% (a.elapsed_time(b) / steps))
''' % ROOT

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="config2")
ap.add_argument("--rounds", type=int, default=3)
ap.add_argument("--steps", type=int, default=8)
ap.add_argument("--pre", type=int, default=3, help="untimed 100-sub-step steps before the timed ones (40: wheels in the groove for config 2)")
ap.add_argument("names", nargs="+")
args = ap.parse_args()
res = {n: [] for n in args.names}
for r in range(args.rounds):
    for n in args.names:
        lib = os.path.join(ROOT, "build", "lib%s.so" % n)
        out = subprocess.run([sys.executable, "-c", ONE, lib, args.workload, str(args.steps), str(args.pre)], capture_output=True, text=True, timeout=600)
        if out.returncode != 0:
            print(n, "FAILED", out.stderr[-400:]); res[n].append(float("nan")); continue
        res[n].append(float(out.stdout.strip().splitlines()[-1]))
for n in args.names:
    print("%-12s %s  best %.3f ms/step" % (n, " ".join("%.3f" % v for v in res[n]), min(res[n])))
