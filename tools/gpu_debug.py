"""First-light GPU vs oracle<float> comparison (developer tool; prints max diffs per step)."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

import oracle_lib  # noqa: E402
from tf_ode_b200 import RoverBatch, abi, glorot_policy, make_config, synthetic_heightmap  # noqa: E402


def bodies_of(s):
    return oracle_lib.state_bodies(s)


def compare(tag, cfg_fn, steps, every=10, prec="float"):
    cfg_g = cfg_fn()
    cfg_o = cfg_fn()
    g = RoverBatch(cfg_g)
    o = oracle_lib.Oracle(cfg_o, prec)
    if cfg_g.controller == abi.CTRL_MLP:
        w = glorot_policy()
        g.load_policy(w)
        o.load_policy(w)
    worst = 0.0
    for k in range(steps):
        g.step(1)
        o.step(1)
        if k % every == every - 1 or k < 3:
            sg, so = g.get_state(0), o.get_state(0)
            bg, bo = bodies_of(sg), bodies_of(so)
            nb = 12 if cfg_g.scene == abi.SCENE_STEP_BOXES else 10
            dp = np.abs(bg[:nb, :7] - bo[:nb, :7]).max()
            dv = np.abs(bg[:nb, 7:] - bo[:nb, 7:]).max()
            worst = max(worst, dp)
            print("%s step %4d dpos %.2e dvel %.2e nc %d/%d seed %08x/%08x ft0 %.3f/%.3f x %.4f/%.4f act %.2f/%.2f t %d/%d ep %d/%d" % (
                tag, k + 1, dp, dv, sg.n_contacts, so.n_contacts, sg.lcg_seed, so.lcg_seed, sg.ft[2], so.ft[2],
                bg[0, 0], bo[0, 0], sg.steering_rate, so.steering_rate, sg.time_steps, so.time_steps, sg.episodes_done, so.episodes_done))
            sys.stdout.flush()
    g.close()
    o.close()
    return worst


if __name__ == "__main__":
    print(torch.cuda.get_device_name(0))
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which in ("all", "plane"):
        compare("plane/ext", lambda: make_config(abi.PRESET_SCENE_1, 1, scene=abi.SCENE_PLANE), 300, 25)
    if which in ("all", "crawl"):
        compare("crawler", lambda: make_config(abi.PRESET_CRAWLING, 1), 300, 25)
    if which in ("all", "steps"):
        compare("steps/random", lambda: make_config(abi.PRESET_MT_RANDOM, 1), 300, 25)
    if which in ("all", "hf"):
        hm = synthetic_heightmap("rock_groove")
        compare("hf/random", lambda: make_config(abi.PRESET_MT_RANDOM, 1, heightfield=(hm, 3.0, 3.0, (1.0, 0.0, -0.3))), 300, 25)
    if which in ("all", "mlp"):
        compare("train/mlp", lambda: make_config(abi.PRESET_TRAINING, 1), 1600, 100)
    if which in ("all", "perf"):
        for n in (4096, 16384):
            hm = synthetic_heightmap("rock_groove")
            cfg = make_config(abi.PRESET_MT_RANDOM, n, heightfield=(hm, 3.0, 3.0, (1.0, 0.0, -0.3)))
            g = RoverBatch(cfg)
            g.step(100)
            g.sync()
            t = time.time()
            g.step(500)
            g.sync()
            dt = time.time() - t
            print("perf n=%d: %.3f s for 500 steps -> %.3e env-steps/s" % (n, dt, n * 500 / dt))
            g.close()
