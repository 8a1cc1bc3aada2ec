#!/usr/bin/env python
"""Per-config histogram of contacts GENERATED per world step (before the engine's per-environment capacity), from the CPU
oracle (test infrastructure) over whole episodes.  Usage: python tools/contact_histogram.py [envs] [steps] [cap]
The reference creates every contact (ode/environment.cc:101-128: 10 per pair, no per-world cap); the engine's capacity must
therefore sit above the largest count any benchmarked config reaches."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
import oracle_lib  # noqa: E402
from tf_ode_b200 import abi, glorot_policy  # noqa: E402

envs = int(sys.argv[1]) if len(sys.argv) > 1 else 256
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 60000
cap = int(sys.argv[3]) if len(sys.argv) > 3 else 0
only = sys.argv[4:] or ["config2", "config3", "config4", "config5"]
for wl in only:
    cfg = bench.make_cfg(wl, envs)
    if cap:
        cfg.max_contacts = cap
    o = oracle_lib.Oracle(cfg, "float", threads=os.cpu_count())
    if cfg.controller == abi.CTRL_MLP:
        o.load_policy(glorot_policy())
    o.step(steps)
    h = o.contact_hist()
    tot = int(h.sum())
    nz = np.nonzero(h)[0]
    eps = sum(o.get_state(e).episodes_done for e in range(envs))
    print("%s: %d envs x %d steps (cap %d), %d episodes ended; max generated = %d" % (wl, envs, steps, cfg.max_contacts, eps, nz.max()))
    cum = 0
    for k in nz:
        cum += int(h[k])
        print("   n=%3d  steps=%10d  (%.4f %%)  cum %.4f %%" % (k, int(h[k]), 100.0 * int(h[k]) / tot, 100.0 * cum / tot))
    for c in (24, 32, 40, 48, 64):
        print("   steps with more than %d generated contacts: %d" % (c, int(h[c + 1:].sum())))
    o.close()
