"""Scene presets: the constants of the reference's main() scenes as BatchConfig values.

Mirrors tfode_batch_config_default() in csrc/tfode_capi.cu (tests check they agree).
  PRESET_SCENE_1    src/scene_1.cc:19-29,68-96        external commands, IC off, never terminates
  PRESET_MT_RANDOM  src/scene_1_mt.cc:20-35,72-132    random actions replace the model tree (BASELINE config 2)
  PRESET_TRAINING   src/rover_training_1.cc:41-146    Rover_1_tf, step boxes, e-greedy, randomised yaw / IC start
  PRESET_CRAWLING   src/crawling.cc:19-27,41-66       Crawler_1 on the bare plane
"""
import ctypes as C
import os

import numpy as np

from . import abi

PRESET_MT_TREE = 4  # Python-side preset: src/scene_1_mt.cc as shipped (Rover_1_mt with two model trees, step boxes)


def make_config(preset, n_envs=1, **kw):
    c = abi.BatchConfig()
    c.abi_version = abi.ABI_VERSION
    c.n_envs = n_envs
    c.device = 0
    c.scene = abi.SCENE_STEP_BOXES
    c.controller = abi.CTRL_EXTERNAL
    c.policy_precision = abi.POLICY_FP32
    c.env_id_offset = 0
    c.auto_reset = 1
    c.mu = 0.5
    c.dt = 0.001
    c.erp = 0.2
    c.cfm = 1e-10
    c.pgs_iters = 20
    c.sor_w = 1.3
    c.crawling_mode = 1
    c.ic_period = 0.0
    c.ic_initially_active = 0
    c.exploration = 0
    c.fork_callback = 0
    c.speedf = 0.04
    c.term = 0.5
    c.ic_start = -1.0
    c.ic_start_jitter = 0.0
    c.timeout = 60.0
    c.x_goal = 2.0
    c.y_max = 0.5
    c.x_goal_abs = 1
    c.check_upside_down = 1
    c.terminate = 0
    c.step_yaw_deg = 0.0
    c.step_yaw_jitter_deg = 0.0
    c.crawler_amp, c.crawler_dt_torque, c.crawler_rate = 30.0, 2.0, 10.0
    c.crawler_amp_jitter = c.crawler_rate_jitter = 0.0
    c.hf_ncol = c.hf_nrow = 0
    c.hf_length = c.hf_width = 3.0
    c.hf_pos[0] = c.hf_pos[1] = c.hf_pos[2] = 0.0
    c.seed = 20260411
    c.max_contacts = 24
    c.transition_capacity = 0
    c.solver = abi.SOLVER_QUICKSTEP
    c.reserved_ = 0
    if preset == abi.PRESET_SCENE_1:
        pass
    elif preset == abi.PRESET_MT_RANDOM:
        c.controller = abi.CTRL_RANDOM
        c.ic_period = 0.5
        c.ic_start = 1.0
        c.x_goal, c.y_max, c.terminate = 1.0, 0.6, 1
    elif preset == PRESET_MT_TREE:  # src/scene_1_mt.cc:20-35,54-62,72-132
        c.controller = abi.CTRL_MODEL_TREE
        c.ic_period = 0.5
        c.ic_start = 1.0
        c.x_goal, c.y_max, c.terminate = 1.0, 0.6, 1
    elif preset == abi.PRESET_TRAINING:
        c.controller = abi.CTRL_MLP
        c.ic_period = 0.5
        c.ic_start, c.ic_start_jitter = 1.0, 0.25
        c.exploration, c.fork_callback = 1, 1
        c.x_goal, c.y_max, c.terminate = 1.5, 0.6, 1
        c.step_yaw_jitter_deg = 5.0
    elif preset == abi.PRESET_CRAWLING:
        c.scene = abi.SCENE_PLANE
        c.controller = abi.CTRL_CRAWLER
        c.mu = 0.6
        c.ic_initially_active = 1
        c.term = 0.0
        c.x_goal, c.y_max, c.x_goal_abs, c.check_upside_down, c.terminate = 10.0, 0.0, 0, 0, 1
    else:
        raise ValueError("unknown preset %r" % (preset,))
    keep = None
    for k, v in kw.items():
        if k == "heightfield":
            keep = set_heightfield(c, *v)
        elif k == "hf_pos":
            c.hf_pos[0], c.hf_pos[1], c.hf_pos[2] = v
        else:
            if not hasattr(c, k):
                raise AttributeError(k)
            setattr(c, k, v)
    c._keepalive = keep
    return c


def set_heightfield(cfg, heights, length=3.0, width=3.0, pos=(0.0, 0.0, 0.0)):
    """heights: [nrow, ncol] float32 metres, row 0 = image top row (ode/heightfield.hh:50-54 after the row flip)."""
    h = np.ascontiguousarray(heights, dtype=np.float32)
    cfg.scene = abi.SCENE_HEIGHTFIELD
    cfg.hf_nrow, cfg.hf_ncol = h.shape
    cfg.hf_length, cfg.hf_width = length, width
    cfg.hf_pos[0], cfg.hf_pos[1], cfg.hf_pos[2] = pos
    cfg.hf_heights = h.ctypes.data_as(C.POINTER(C.c_float))
    cfg._keepalive = h
    return h


_LEVELS = None


def reference_heightmap(kind):
    """The reference's own terrain: red-channel levels of env_data/heightmap_<kind>.png (committed as
    tf_ode_b200/data/heightmap_levels.npz -- package data, written by tools/make_golden.py, since /root/reference does not travel) turned into
    heights exactly as ode/heightfield.hh:54 does (level * 0.3 / 255 in double), float32 [nrow, ncol], row 0 = top row.
    kind: 'rock_step' | 'rock_groove' | 'rock_groove_large'."""
    global _LEVELS
    if _LEVELS is None:
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "heightmap_levels.npz")
        _LEVELS = dict(np.load(path))
    return (_LEVELS[kind].astype(np.float64) * 0.3 / 255).astype(np.float32)


def synthetic_heightmap(kind, n=200):
    """Synthetic stand-ins with the statistics of env_data/heightmap_rock_{step,groove}.png (the PNGs stay in
    the reference tree; there is no /root/reference on the GPU box).  Values are 8-bit levels * 0.3/255 like
    ode/heightfield.hh:54.  kind: 'rock_step' | 'rock_groove' | 'rock_groove_large'."""
    if kind == "rock_groove_large":
        n = 400
    rng = np.random.RandomState(7 if kind == "rock_step" else 11)
    yy, xx = np.mgrid[0:n, 0:n].astype(np.float64) / (n - 1)
    bump = np.zeros((n, n))
    for _ in range(40):
        cx, cy, s, a = rng.rand(), rng.rand(), 0.02 + 0.05 * rng.rand(), rng.randn()
        bump += a * np.exp(-((xx - cx) ** 2 + (yy - cy) ** 2) / (2 * s * s))
    bump /= np.abs(bump).max()
    if kind == "rock_step":
        # a rock plateau ~0.27 m high in the middle of a flat zero field, rough top, ragged edge
        r = np.sqrt(((xx - 0.42) / 0.30) ** 2 + ((yy - 0.5) / 0.38) ** 2) + 0.08 * bump
        lvl = np.clip((1.0 - r) * 6.0, 0.0, 1.0) * (215 + 35 * bump)
    else:
        # flat plateau at 255 with a meandering groove down to ~90 crossing the map top to bottom
        scale = 1.0 if n == 200 else 0.5
        centre = 0.45 + 0.12 * yy + 0.03 * np.sin(yy * 9.0)
        wdt = (0.035 + 0.01 * np.sin(yy * 5.0)) * scale
        d = np.abs(xx - centre) / wdt
        lvl = 255 - np.clip(1.6 - d, 0.0, 1.0) * (165 - 10 * bump)
        lvl[(yy < 0.05) | (yy > 0.97)] = 255
    lvl = np.clip(np.round(lvl), 0, 255)
    return (lvl * (0.3 / 255.0)).astype(np.float32)
