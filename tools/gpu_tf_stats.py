"""Teacher-forced single-step GPU vs oracle<float> statistics (developer tool used to set the test tolerances)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib  # noqa: E402
from tf_ode_b200 import RoverBatch, abi, glorot_policy, make_config, synthetic_heightmap  # noqa: E402


def run(tag, cfg_fn, steps):
    g = RoverBatch(cfg_fn()); o = oracle_lib.Oracle(cfg_fn(), "float")
    if g.cfg.controller == abi.CTRL_MLP:
        w = glorot_policy(); g.load_policy(w); o.load_policy(w)
    nb = 12 if g.cfg.scene == abi.SCENE_STEP_BOXES else 10
    mx = np.zeros(4); nc_bad = 0; seed_bad = 0; other = 0; worst_k = -1
    for k in range(steps):
        s = o.get_state(0)
        if s.episode != g.get_state(0).episode:
            g.reset([0])
        g.set_state(0, s)
        g.step(1); o.step(1)
        sg, so = g.get_state(0), o.get_state(0)
        bg, bo = oracle_lib.state_bodies(sg)[:nb], oracle_lib.state_bodies(so)[:nb]
        d = np.array([np.abs(bg[:, 0:3] - bo[:, 0:3]).max(), np.abs(bg[:, 3:7] - bo[:, 3:7]).max(),
                      np.abs(bg[:, 7:10] - bo[:, 7:10]).max(), np.abs(bg[:, 10:13] - bo[:, 10:13]).max()])
        if sg.n_contacts != so.n_contacts:
            nc_bad += 1
            continue
        if d[3] > mx[3]:
            worst_k = k
        mx = np.maximum(mx, d)
        seed_bad += sg.lcg_seed != so.lcg_seed
        for f in ("tick", "time_steps", "episodes_done", "ic_activated", "term_reason", "crawler_phase", "explore"):
            other += getattr(sg, f) != getattr(so, f)
        fa = np.abs(np.array(sg.ft) - np.array(so.ft)).max()
        mx_ft = fa
    print("%-14s steps %d  dpos %.2e dquat %.2e dlvel %.2e davel %.2e (worst step %d)  nc mismatches %d  seed mismatches %d  other %d  ft %.2e" % (
        tag, steps, mx[0], mx[1], mx[2], mx[3], worst_k, nc_bad, seed_bad, other, mx_ft))
    sys.stdout.flush()
    g.close(); o.close()


if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 3000
    hm = synthetic_heightmap("rock_groove")
    hs = synthetic_heightmap("rock_step")
    run("plane/ext", lambda: make_config(abi.PRESET_SCENE_1, 1, scene=abi.SCENE_PLANE), n)
    run("crawler", lambda: make_config(abi.PRESET_CRAWLING, 1), n)
    run("steps/random", lambda: make_config(abi.PRESET_MT_RANDOM, 1), n)
    run("hf_groove/rnd", lambda: make_config(abi.PRESET_MT_RANDOM, 1, heightfield=(hm, 3.0, 3.0, (1.0, 0.0, -0.3))), n)
    run("hf_step/ext", lambda: make_config(abi.PRESET_SCENE_1, 1, heightfield=(hs, 3.0, 3.0, (2.0, 0.0, -0.01))), n)
    run("train/mlp", lambda: make_config(abi.PRESET_TRAINING, 1), n)
    run("train/fast", lambda: make_config(abi.PRESET_TRAINING, 1, speedf=0.5, x_goal=0.6), 2 * n)
