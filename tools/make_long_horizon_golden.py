#!/usr/bin/env python
"""Writes tests/golden/long_horizon_<scene>.npz: one complete episode per environment of oracle<double> (the reference-equivalent fp64
restatement) for the two scenes of the long-horizon contract -- per-episode return, length, termination reason, final position.

  python tools/make_long_horizon_golden.py            (CPU only; about 10 minutes on 8 cores)

The GPU test (tests/test_gpu_configs.py::test_long_horizon_*) compares the CUDA path's episode statistics with these arrays, so that the GPU
box does not spend six minutes of host time re-running the oracle; TFODE_LONG_HORIZON_LIVE=1 makes it run the oracle live again (and check it
against this file).  tests/test_oracle_kat.py re-runs the first environments of each scene on the CPU and requires them equal to the file:
the per-episode random draws are keyed on the global environment id, so a 32-environment oracle reproduces rows 0..31 exactly.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

N_EPISODES = 4096
SUBSTEPS = 62000   # 60 s timeout + the latest IC start


def scene_config(scene, n):
    """The two scenes of the contract; the GPU test builds its batch from the same function."""
    import bench
    from tf_ode_b200 import abi, make_config
    if scene == "training_fast":
        return make_config(abi.PRESET_TRAINING, n, speedf=0.5, x_goal=0.6, auto_reset=0)
    if scene == "config2":
        return bench.make_cfg("config2", n, auto_reset=0)
    raise ValueError(scene)


def oracle_episodes(scene, n, threads):
    import oracle_lib
    from tf_ode_b200 import abi, glorot_policy
    cfg = scene_config(scene, n)
    o = oracle_lib.Oracle(cfg, "double", threads=threads)
    if cfg.controller == abi.CTRL_MLP:
        o.load_policy(glorot_policy())
    o.step(SUBSTEPS)
    s = [o.get_state(e) for e in range(n)]
    o.close()
    return dict(total_reward=np.array([x.total_reward for x in s], np.float64), time_steps=np.array([x.time_steps for x in s], np.int64),
                term_reason=np.array([x.term_reason for x in s], np.int32), episodes_done=np.array([x.episodes_done for x in s], np.int32),
                pos_xy=np.array([[x.pos[0][0], x.pos[0][1]] for x in s], np.float64))


def main():
    for scene in ("training_fast", "config2"):
        d = oracle_episodes(scene, N_EPISODES, os.cpu_count() or 8)
        assert (d["episodes_done"] == 1).all()
        path = os.path.join(ROOT, "tests", "golden", "long_horizon_%s.npz" % scene)
        np.savez_compressed(path, n=N_EPISODES, substeps=SUBSTEPS, **d)
        print(path, "return mean %.6g, length mean %.6g" % (d["total_reward"].mean(), d["time_steps"].mean()))


if __name__ == "__main__":
    main()
