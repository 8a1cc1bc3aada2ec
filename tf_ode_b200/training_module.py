"""Batched stand-in for the reference's Boost.Python trainer module `rover_training_1_module`
(src/rover_training_1.cc:243-261): `trial()` and `eval()` over many environments at once.

Reference:  trial(path_to_model_dir) -> list of (state[17], (steering_rate, boggie_torque), reward, terminal, next_state[17])
            for ONE episode of ONE rover driven by the SavedModel actor at that path with e-greedy exploration
            (simulation("trial", ...), src/rover_training_1.cc:41-224); eval(path) runs one episode without exploration
            and prints the result line (:203-209).
Here:       the actor is either that same SavedModel directory (a str / os.PathLike: its checkpoint variables are read by
            tfode_savedmodel_read_actor, no TensorFlow involved) or the fp32 weight blob the engine takes
            (tfode_batch_load_policy), and one call runs `n_episodes` independent episodes in one batch on the GPU.  trial() returns one such list
            per episode, built from the engine's transition ring; the terminal patch and the end-of-episode penalties
            (-2 beyond y_max, -5 upside-down, :212-224) are applied on the device.
There is no CPU path: this calls the C-ABI library and needs a B200.
"""
import os

import numpy as np

from . import abi
from .config import make_config
from .engine import RoverBatch, read_saved_model_actor

TIMEOUT_STEPS = 60000  # float timeout(60) at 1 ms steps (src/rover_training_1.cc:119)


def _actor(actor):
    """path_to_model_dir (the reference's argument) or a weight blob -> float32 blob."""
    if isinstance(actor, (str, bytes, os.PathLike)):
        return read_saved_model_actor(actor)
    return np.asarray(actor, np.float32)


def _run_episodes(actor_weights, n_episodes, exploration, seed, device, policy_precision, cfg_overrides):
    import torch
    kw = dict(auto_reset=0, exploration=int(exploration), seed=seed, device=device, policy_precision=policy_precision,
              transition_capacity=n_episodes * (TIMEOUT_STEPS // 500 + 2))
    kw.update(cfg_overrides or {})
    cfg = make_config(abi.PRESET_TRAINING, n_episodes, **kw)
    batch = RoverBatch(cfg, stream=torch.cuda.current_stream().cuda_stream)
    try:
        batch.load_policy(actor_weights)
        done = torch.zeros(n_episodes, dtype=torch.uint8, device="cuda")
        finished = torch.zeros(n_episodes, dtype=torch.bool, device="cuda")
        steps = 0
        while steps < TIMEOUT_STEPS + 1000:
            batch.step(1000)
            steps += 1000
            batch.observe(done=done)
            finished |= done.bool()
            if bool(finished.all()):
                break
        cap = cfg.transition_capacity
        buf = torch.zeros(cap, abi.TRANSITION_WIDTH, device="cuda")
        n = batch.drain_transitions(buf, cap)
        if batch.last_dropped:
            raise RuntimeError("%d experience rows were lost: raise transition_capacity" % batch.last_dropped)
        rows = buf[:n].cpu().numpy()
        states = [batch.get_state(e) for e in range(n_episodes)]
    finally:
        batch.close()
    return rows, states, cfg


def rows_to_experience(rows, n_episodes):
    """Transition rows (env id, s[17], a[2], r, terminal, s'[17]) -> per-episode lists of the reference's 5-tuples."""
    out = [[] for _ in range(n_episodes)]
    for r in rows:
        out[int(r[0])].append((r[1:18].tolist(), (float(r[18]), float(r[19])), float(r[20]), bool(r[21] != 0), r[22:39].tolist()))
    return out


def trial(actor_weights, n_episodes=1, seed=20260411, device=0, policy_precision=abi.POLICY_FP32, **cfg_overrides):
    """`n_episodes` x simulation("trial", actor): exploration on, randomised step yaw and IC start.  Returns a list of
    `n_episodes` experience lists; with n_episodes == 1 the single list itself, like the reference's trial()."""
    rows, _, _ = _run_episodes(_actor(actor_weights), n_episodes, True, seed, device, policy_precision, cfg_overrides)
    exp = rows_to_experience(rows, n_episodes)
    return exp[0] if n_episodes == 1 else exp


def result_line(ok, t, x, y, rmoy, colour=True):
    """The line simulation() prints (src/rover_training_1.cc:205-207), parsed by scripts/bin/monitor-policies."""
    tag = "[Success]" if ok else "[Failure]"
    if colour:
        tag = ("\033[1;32m" if ok else "\033[1;31m") + tag + "\033[0;39m"
    return "%s t %6.3f | x %5.3f | y %+6.3f | Rmoy %7.3f" % (tag, t, x, y, rmoy)


def eval(actor_weights, n_episodes=1, seed=20260411, device=0, policy_precision=abi.POLICY_FP32, colour=True, file=None, **cfg_overrides):  # noqa: A001 (the reference's name)
    """`n_episodes` x simulation("eval", actor): no exploration; prints one result line per episode and returns
    (success, t, x, y, Rmoy) tuples."""
    _, states, cfg = _run_episodes(_actor(actor_weights), n_episodes, False, seed, device, policy_precision, cfg_overrides)
    getpos = np.array([-0.1095, -0.023, 0.0075])  # Rover_1::GetPosition's point of the main body (src/rover_1.cc:248-253)
    res = []
    for s in states:
        w, x_, y_, z_ = s.quat[0]
        R = np.array([[1 - 2 * (y_ * y_ + z_ * z_), 2 * (x_ * y_ - w * z_), 2 * (x_ * z_ + w * y_)],
                      [2 * (x_ * y_ + w * z_), 1 - 2 * (x_ * x_ + z_ * z_), 2 * (y_ * z_ - w * x_)],
                      [2 * (x_ * z_ - w * y_), 2 * (y_ * z_ + w * x_), 1 - 2 * (x_ * x_ + y_ * y_)]])
        p = R @ getpos + np.array(s.pos[0])
        t = float(np.float32(s.time_steps) * np.float32(0.001))  # Sim_loop::get_time (renderer/sim_loop.hh:57)
        ok = abs(p[0]) >= cfg.x_goal
        rmoy = s.total_reward / t if t > 0 else 0.0
        res.append((bool(ok), t, float(p[0]), float(p[1]), float(rmoy)))
        print(result_line(ok, t, p[0], p[1], rmoy, colour), file=file, flush=True)
    return res
