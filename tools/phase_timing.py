"""Developer tool: per-phase cycle split of step_kernel (build with -DTFODE_PHASE_TIMING into a scratch library).

  python tools/phase_timing.py [workload] [envs]      (on the GPU box)
"""
import ctypes as C
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
src = os.path.join(ROOT, "tf_ode_b200", "csrc")
lib = os.path.join(ROOT, "gpurun_out", "libtfode_phase.so")
os.makedirs(os.path.dirname(lib), exist_ok=True)
subprocess.check_call(["/usr/local/cuda/bin/nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-fmad=false", "-DTFODE_PHASE_TIMING", *os.environ.get("TFODE_EXTRA", "").split(),
                       "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off", "-shared", "-cudart", "static", "-o", lib, os.path.join(src, "tfode_capi.cu")])
import tf_ode_b200.engine as eng  # noqa: E402
eng.LIB_PATH = lib
import bench  # noqa: E402
from tf_ode_b200 import RoverBatch, abi, glorot_policy  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "config2"
n = int(sys.argv[2]) if len(sys.argv) > 2 else bench.dev_envs(wl)
g = RoverBatch(bench.make_cfg(wl, n))
if g.cfg.controller == abi.CTRL_MLP:
    g.load_policy(glorot_policy())
out = (C.c_ulonglong * 16)()
g.step(300); g.sync()
g.L.tfode_debug_phase_cycles(out, 1)
g.step(300); g.sync()
g.L.tfode_debug_phase_cycles(out, 0)
names = ["collide", "rows+rhs+order", "wait(before pgs)", "pgs (warp 0)", "wait(after pgs)", "integrate+robot+control"]
tot = float(sum(out[:6]))
ncta = (n + 7) // 8
for nm, v in zip(names, out[:6]):
    print("%-26s %5.1f%%   %8.0f cycles per CTA-substep" % (nm, 100 * v / tot, v / ncta / 300))
print("pgs: %.1f lock-steps per sweep (warp 0 of each CTA), %d sweeps" % (out[6] / max(out[7], 1), out[7]))
for nm, v in zip(["pgs: queue sizes", "pgs: Fisher-Yates", "pgs: queue fill", "pgs: sweeps"], out[8:12]):
    print("  %-24s %8.0f cycles per CTA-substep" % (nm, v / ncta / 300))
for nm, v in zip(["collide: non-sphere geoms vs box/hf", "collide: sphere range tests (hf)", "collide: sphere-hf warp scans", "collide: rest (plane, compaction)"], out[12:16]):
    print("  %-36s %8.0f cycles per CTA-substep" % (nm, v / ncta / 300))
print("  per lock-step: %.0f cycles" % (out[11] / max(out[6], 1) * 1.0))
print("total %.0f cycles per CTA-substep (%s, %d envs)" % (tot / ncta / 300, wl, n))
