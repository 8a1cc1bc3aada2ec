"""Developer tool: locate the first sub-step at which the GPU and oracle<float> differ on config 4 (crawler), with and without a reset."""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import bench, oracle_lib
from tf_ode_b200 import RoverBatch

FIELDS = ("lcg_seed", "n_contacts", "tick", "time_steps", "episode", "episodes_done", "crawler_phase", "servo_mode_vel")
SC = ("servo_vel", "robot_speed", "steering_rate", "boggie_torque", "ic_clock", "crawler_amp", "crawler_rate", "crawler_dtorque_dt", "speed")

def diff(g, o, n):
    bad = []
    for e in range(n):
        sg, so = g.get_state(e), o.get_state(e)
        bg, bo = oracle_lib.state_bodies(sg)[:10], oracle_lib.state_bodies(so)[:10]
        if not np.array_equal(bg, bo) or any(getattr(sg, f) != getattr(so, f) for f in FIELDS):
            bad.append((e, float(np.abs(bg - bo).max()), {f: (getattr(sg, f), getattr(so, f)) for f in FIELDS if getattr(sg, f) != getattr(so, f)},
                        {f: (getattr(sg, f), getattr(so, f)) for f in SC if getattr(sg, f) != getattr(so, f)}))
    return bad

n = 64
first_id = int(sys.argv[1]) if len(sys.argv) > 1 else 512
for mode in ("noreset", "reset"):
    g = RoverBatch(bench.make_cfg("config4", n, env_id_offset=first_id)); o = oracle_lib.Oracle(bench.make_cfg("config4", n, env_id_offset=first_id), "float", threads=16)
    if mode == "reset":
        a = bench.action_tape(0, 0, n, first_id)
        g.set_actions(torch.from_numpy(a).cuda()); o.set_actions(a); g.step(1500); o.step(1500); g.reset(); o.reset()
        print(mode, "after reset:", diff(g, o, n)[:2])
    k = 0
    while k < 12000:
        g.step(250); o.step(250); k += 250
        b = diff(g, o, n)
        if b:
            print(mode, "first difference within sub-steps %d..%d: %d envs" % (k - 250, k, len(b)), b[:3])
            break
    else:
        print(mode, "no difference in 12000 sub-steps")
    g.close(); o.close()
