"""SURVEY 8(f) n1 on the GPU: TFODE_SOLVER_DANTZIG (the dWorldStep equivalent, ode/environment.hh:132) through the C ABI against the
CPU oracle running the same include/tfode_lcp.h: bit-exact in float, within 1e-3 m of oracle<double> over 1000 steps, no stalled
pivoting drive, no random numbers drawn."""
import numpy as np
import pytest

import oracle_lib
from tf_ode_b200 import RoverBatch, abi, glorot_policy, make_config, reference_heightmap

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

SCENES = {
    "plane_ext": lambda n, **kw: make_config(abi.PRESET_SCENE_1, n, scene=abi.SCENE_PLANE, solver=abi.SOLVER_DANTZIG, **kw),
    "steps_random": lambda n, **kw: make_config(abi.PRESET_MT_RANDOM, n, solver=abi.SOLVER_DANTZIG, **kw),
    "training_fast": lambda n, **kw: make_config(abi.PRESET_TRAINING, n, speedf=0.5, x_goal=0.6, solver=abi.SOLVER_DANTZIG, **kw),
    "hf_groove": lambda n, **kw: make_config(abi.PRESET_MT_RANDOM, n, solver=abi.SOLVER_DANTZIG,
                                             heightfield=(reference_heightmap("rock_groove"), 3.0, 3.0, (1.0, 0.0, -0.3)), **kw),
}


@pytest.mark.parametrize("name,steps", [("plane_ext", 1500), ("steps_random", 2500), ("training_fast", 2500), ("hf_groove", 1500)])
def test_dantzig_trajectories_are_bit_exact_against_the_oracle(name, steps):
    n = 16
    g = RoverBatch(SCENES[name](n)); o = oracle_lib.Oracle(SCENES[name](n), "float", threads=8)
    if g.cfg.controller == abi.CTRL_MLP:
        w = glorot_policy(); g.load_policy(w); o.load_policy(w)
    nb = 12 if g.cfg.scene == abi.SCENE_STEP_BOXES else 10
    nc = torch.zeros(n, dtype=torch.int32, device="cuda")
    for k in range(steps // 250):
        g.step(250); o.step(250)
        g.observe(ncontacts=nc)
        assert np.array_equal(nc.cpu().numpy(), o.observe()[3]), "%s step %d: contact counts" % (name, (k + 1) * 250)
    for e in range(n):
        sg, so = g.get_state(e), o.get_state(e)
        bg, bo = oracle_lib.state_bodies(sg)[:nb], oracle_lib.state_bodies(so)[:nb]
        assert np.array_equal(bg, bo), "%s env %d: max |d| = %.3e" % (name, e, np.abs(bg - bo).max())
        assert sg.lcg_seed == so.lcg_seed == 0 and sg.time_steps == so.time_steps and sg.episodes_done == so.episodes_done
    # stalled pivoting drives (lcp.cpp's "s <= 0" exit): the same islands stall on both sides, and binary32 stalls on at most one island
    # solve in a thousand (none with the double oracle, tests/test_lcp.py)
    fails = g.lcp_failures()
    assert fails == oracle_lib.lib().orc_lcp_failures(o.h) and fails <= n * steps * 3 // 1000
    g.close(); o.close()


def test_dantzig_against_double_oracle_and_against_quickstep():
    g = RoverBatch(SCENES["plane_ext"](1)); o = oracle_lib.Oracle(SCENES["plane_ext"](1), "double")
    q = RoverBatch(make_config(abi.PRESET_SCENE_1, 1, scene=abi.SCENE_PLANE))
    g.step(1000); o.step(1000); q.step(1000)
    bg, bo, bq = oracle_lib.state_bodies(g.get_state(0))[:10], o.bodies(0)[:10], oracle_lib.state_bodies(q.get_state(0))[:10]
    assert np.abs(bg[:, 0:3] - bo[:, 0:3]).max() <= 1e-3 and np.abs(bg[:, 3:7] - bo[:, 3:7]).max() <= 5e-3   # fp32 vs fp64, same solver
    assert np.abs(bg[:, 0:3] - bq[:, 0:3]).max() <= 5e-3                                                      # exact LCP vs 20 SOR sweeps
    for h in (g, o, q):
        h.close()
