"""Developer tool: first differing fields between the GPU and oracle<float> (free running)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib
from tf_ode_b200 import RoverBatch, abi, make_config
scene = sys.argv[1] if len(sys.argv) > 1 else "plane"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1
mk = {"plane": lambda: make_config(abi.PRESET_SCENE_1, n, scene=abi.SCENE_PLANE),
      "steps": lambda: make_config(abi.PRESET_MT_RANDOM, n)}[scene]
g = RoverBatch(mk()); o = oracle_lib.Oracle(mk(), "float")
names = ["px","py","pz","qw","qx","qy","qz","vx","vy","vz","wx","wy","wz"]
for k in range(60):
    g.step(1); o.step(1)
    bad = False
    for e in range(n):
        sg, so = g.get_state(e), o.get_state(e)
        bg, bo = oracle_lib.state_bodies(sg), oracle_lib.state_bodies(so)
        d = np.argwhere(bg != bo)
        if len(d) or sg.lcg_seed != so.lcg_seed:
            print("step", k + 1, "env", e, "seed", hex(sg.lcg_seed), hex(so.lcg_seed), "nc", sg.n_contacts, so.n_contacts)
            for b, f in d[:12]:
                print("  body", b, names[f], repr(bg[b, f]), repr(bo[b, f]))
            bad = True
    if bad:
        break
else:
    print("no difference in 60 steps")
