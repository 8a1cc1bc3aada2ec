"""GPU (CUDA path through the C ABI) vs the CPU oracle.  Run with `-m gpu` on a B200."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib
from tf_ode_b200 import RoverBatch, abi, glorot_policy, make_config, synthetic_heightmap

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

HM_GROOVE = synthetic_heightmap("rock_groove")
HM_STEP = synthetic_heightmap("rock_step")

SCENES = {
    "plane_ext": lambda n=1, **kw: make_config(abi.PRESET_SCENE_1, n, scene=abi.SCENE_PLANE, **kw),
    "crawler": lambda n=1, **kw: make_config(abi.PRESET_CRAWLING, n, crawler_amp=25.0, crawler_amp_jitter=5.0, crawler_rate_jitter=5.0, **kw),
    "steps_random": lambda n=1, **kw: make_config(abi.PRESET_MT_RANDOM, n, **kw),
    "hf_groove_random": lambda n=1, **kw: make_config(abi.PRESET_MT_RANDOM, n, heightfield=(HM_GROOVE, 3.0, 3.0, (1.0, 0.0, -0.3)), **kw),
    "hf_rock_step_ext": lambda n=1, **kw: make_config(abi.PRESET_SCENE_1, n, speedf=0.5, heightfield=(HM_STEP, 3.0, 3.0, (2.0, 0.0, -0.01)), **kw),
    "training_mlp": lambda n=1, **kw: make_config(abi.PRESET_TRAINING, n, **kw),
    "training_fast": lambda n=1, **kw: make_config(abi.PRESET_TRAINING, n, speedf=0.5, x_goal=0.6, **kw),
}


def make_pair(name, n=1, prec="float", **kw):
    g = RoverBatch(SCENES[name](n, **kw))
    o = oracle_lib.Oracle(SCENES[name](n, **kw), prec)
    if g.cfg.controller == abi.CTRL_MLP:
        w = glorot_policy()
        g.load_policy(w)
        o.load_policy(w)
    return g, o


def nbodies(cfg):
    return 12 if cfg.scene == abi.SCENE_STEP_BOXES else 10


INT_FIELDS = ("lcg_seed", "tick", "time_steps", "episode", "episodes_done", "ic_activated", "term_reason", "crawler_phase",
              "explore", "has_last_state", "fork_collision", "n_contacts", "done")


def assert_states_equal(sg, so, nb, exact=True, tag=""):
    bg, bo = oracle_lib.state_bodies(sg)[:nb], oracle_lib.state_bodies(so)[:nb]
    for f in INT_FIELDS:
        assert getattr(sg, f) == getattr(so, f), "%s: %s %r != %r" % (tag, f, getattr(sg, f), getattr(so, f))
    if exact:
        assert np.array_equal(bg, bo), "%s: body state differs, max |d| = %.3e" % (tag, np.abs(bg - bo).max())
        for f in ("facc", "tacc", "ft", "filt", "wheel_vel", "wheel_fmax", "W", "last_state", "last_pos"):
            a, b = np.array(getattr(sg, f)), np.array(getattr(so, f))
            assert np.array_equal(a[:nb] if a.shape[0] == abi.NBODY else a, b[:nb] if b.shape[0] == abi.NBODY else b), "%s: %s" % (tag, f)
        for f in ("servo_vel", "robot_speed", "steering_rate", "boggie_torque", "ic_clock", "total_reward", "last_reward", "speed"):
            assert getattr(sg, f) == getattr(so, f), "%s: %s %r != %r" % (tag, f, getattr(sg, f), getattr(so, f))


# ----------------------------------------------------------------------------------------------------------------------
# 1. bit-exact free-running trajectories, GPU fp32 vs oracle<float> (strict numerics contract, DESIGN.md "Numerics")
# ----------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name,steps", [("plane_ext", 3000), ("crawler", 3000), ("steps_random", 4000), ("hf_groove_random", 4000),
                                        ("hf_rock_step_ext", 4000), ("training_mlp", 3000), ("training_fast", 6000)])
def test_free_running_trajectory_is_bit_exact(name, steps):
    g, o = make_pair(name, 1)
    nb = nbodies(g.cfg)
    k = 0
    chunk = 1
    while k < steps:
        g.step(chunk); o.step(chunk); k += chunk
        assert_states_equal(g.get_state(0), o.get_state(0), nb, True, "%s step %d" % (name, k))
        chunk = 1 if k < 200 else 37  # every step at first (contact counts bit-exact every step), then strided
    g.close(); o.close()


@pytest.mark.parametrize("name", ["steps_random", "hf_groove_random", "crawler", "training_fast"])
def test_batch_of_envs_matches_oracle_bit_exact_including_contact_counts(name):
    n, steps = 64, 1500
    g, o = make_pair(name, n)
    nc_g = torch.zeros(n, dtype=torch.int32, device="cuda")
    for k in range(steps // 50):
        g.step(50); o.step(50)
        g.observe(ncontacts=nc_g)
        _, _, _, nc_o = o.observe()
        assert np.array_equal(nc_g.cpu().numpy(), nc_o), "%s step %d" % (name, (k + 1) * 50)
    nb = nbodies(g.cfg)
    for e in range(0, n, 7):
        assert_states_equal(g.get_state(e), o.get_state(e), nb, True, "%s env %d" % (name, e))
    g.close(); o.close()


# ----------------------------------------------------------------------------------------------------------------------
# 2. GPU fp32 vs the reference-equivalent fp64 oracle: stated float tolerance over a short horizon
# ----------------------------------------------------------------------------------------------------------------------
def test_short_horizon_against_double_oracle_flat_ground():
    g, o = make_pair("plane_ext", 1, "double")
    g.step(1000); o.step(1000)
    bg, bo = oracle_lib.state_bodies(g.get_state(0))[:10], o.bodies(0)[:10]
    assert np.abs(bg[:, 0:3] - bo[:, 0:3]).max() <= 1e-3   # metres, 1000 steps
    assert np.abs(bg[:, 3:7] - bo[:, 3:7]).max() <= 5e-3   # quaternion components (1e-2 rad)
    g.close(); o.close()


def test_first_steps_against_double_oracle_are_tight():
    for name in ("steps_random", "hf_groove_random"):
        g, o = make_pair(name, 1, "double")
        for k in range(20):  # before the first contacts chatter: pure joint/FT dynamics
            g.step(1); o.step(1)
            sg, so = g.get_state(0), o.get_state(0)
            assert sg.n_contacts == so.n_contacts and sg.lcg_seed == so.lcg_seed
        nb = nbodies(g.cfg)
        bg, bo = oracle_lib.state_bodies(sg)[:nb], oracle_lib.state_bodies(so)[:nb]
        assert np.abs(bg[:, 0:7] - bo[:, 0:7]).max() <= 1e-6 and np.abs(bg[:, 7:13] - bo[:, 7:13]).max() <= 2e-3
        g.close(); o.close()


# (the long-horizon statistics contract -- >= 4096 complete episodes against oracle<double> -- lives in tests/test_gpu_configs.py)


# ----------------------------------------------------------------------------------------------------------------------
# 3. policy op (TF_model::infer) vs a float32 numpy MLP
# ----------------------------------------------------------------------------------------------------------------------
def numpy_mlp(w, x):
    H = 512
    o = 0
    W1 = w[o:o + 17 * H].reshape(17, H); o += 17 * H
    b1 = w[o:o + H]; o += H
    W2 = w[o:o + H * H].reshape(H, H); o += H * H
    b2 = w[o:o + H]; o += H
    W3 = w[o:o + H * 2].reshape(H, 2); o += H * 2
    b3 = w[o:o + 2]
    h = np.maximum(x.astype(np.float64) @ W1 + b1, 0)
    h = np.maximum(h @ W2 + b2, 0)
    return np.tanh(h @ W3 + b3)


@pytest.mark.parametrize("m", [1, 7, 8, 1000, 16384])
def test_policy_infer_matches_numpy(m):
    g = RoverBatch(make_config(abi.PRESET_TRAINING, 8))
    w = glorot_policy()
    w[17 * 512:17 * 512 + 512] = np.random.RandomState(1).randn(512).astype(np.float32) * 0.1  # non-zero biases
    g.load_policy(w)
    x = (np.random.RandomState(m).rand(m, 17).astype(np.float32) * 2 - 1)
    xd = torch.from_numpy(x).cuda()
    yd = torch.zeros(m, 2, device="cuda")
    g.policy_infer(xd, yd)
    torch.cuda.synchronize()
    assert np.abs(yd.cpu().numpy() - numpy_mlp(w, x)).max() <= 2e-5  # fp32 accumulate vs fp64 reference of the same weights
    g.close()


def test_policy_infer_is_bit_exact_with_oracle_fp32_path():
    g, o = make_pair("training_mlp", 4)
    x = (np.random.RandomState(5).rand(333, 17).astype(np.float32) * 2 - 1)
    y = np.zeros((333, 2), np.float32)
    o.L.orc_mlp_infer(o.h, x.ctypes.data_as(C.c_void_p), 333, y.ctypes.data_as(C.c_void_p))
    yd = torch.zeros(333, 2, device="cuda")
    g.policy_infer(torch.from_numpy(x).cuda(), yd)
    assert np.array_equal(yd.cpu().numpy(), y)
    g.close(); o.close()


def test_policy_from_saved_model_directory_equals_the_blob(tmp_path):
    """Rover_1_tf's `actor_path` (src/rover_1_tf.cc:24): the SavedModel directory installs the same actor as its fp32 blob."""
    from test_formats import _actor_tensors, write_bundle
    t, blob = _actor_tensors(seed=11)
    for k in t:
        if t[k][1] == 1:
            t[k][0][...] *= 0.05
    blob = blob * np.float32(0.05)
    write_bundle(str(tmp_path / "actor" / "variables" / "variables"), t, num_shards=2, entries_per_block=4)
    x = torch.from_numpy(np.random.RandomState(2).rand(500, 17).astype(np.float32) * 2 - 1).cuda()
    ya, yb = torch.zeros(500, 2, device="cuda"), torch.zeros(500, 2, device="cuda")
    g = RoverBatch(make_config(abi.PRESET_TRAINING, 8))
    g.load_policy_saved_model(str(tmp_path / "actor"))
    g.policy_infer(x, ya)
    g.load_policy(blob)
    g.policy_infer(x, yb)
    torch.cuda.synchronize()
    assert torch.equal(ya, yb) and float(ya.abs().max()) > 0.01
    g.close()


def _bf16(a):
    """Round fp32 to bf16 (nearest even) and back, the rounding __float2bfloat16_rn applies to the MMA operands."""
    u = np.ascontiguousarray(a, np.float32).view(np.uint32).astype(np.uint64)
    u = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16) << 16
    return u.astype(np.uint32).view(np.float32).reshape(np.shape(a))


def numpy_mlp_bf16(w, x):
    """The tensor-core path's arithmetic: bf16 operands, wide accumulation, fp32 bias/ReLU, bf16 hidden activations
    into layer 2, fp32 output layer (tf_ode_b200/csrc/tfode_policy_tc.cuh)."""
    H = 512
    o = 0
    W1 = w[o:o + 17 * H].reshape(17, H); o += 17 * H
    b1 = w[o:o + H]; o += H
    W2 = w[o:o + H * H].reshape(H, H); o += H * H
    b2 = w[o:o + H]; o += H
    W3 = w[o:o + H * 2].reshape(H, 2); o += H * 2
    b3 = w[o:o + 2]
    h = np.maximum((_bf16(x).astype(np.float64) @ _bf16(W1).astype(np.float64)).astype(np.float32) + b1, 0)
    h = np.maximum((_bf16(h).astype(np.float64) @ _bf16(W2).astype(np.float64)).astype(np.float32) + b2, 0)
    return np.tanh(h.astype(np.float64) @ W3 + b3)


@pytest.mark.parametrize("m", [1, 7, 128, 129, 1000, 16384])
def test_policy_tc_matches_bf16_reference(m):
    """tcgen05 kernel vs numpy with the same operand rounding: the fp32 accumulation order differs (the kernel adds the biases inside
    the tensor core's accumulation), and a hidden activation that lands within an fp32 ulp of a bf16 rounding
    boundary may round the other way (one bf16 ulp of one of 512 activations): <= 5e-4 on a tanh output;
    vs the fp64 MLP on unrounded weights the bf16 operands cost <= 3e-2 on a tanh output (stated tolerance of the
    TFODE_POLICY_TC path; the fp32 kernel stays the bit-exact parity path)."""
    g = RoverBatch(make_config(abi.PRESET_TRAINING, 8, policy_precision=abi.POLICY_TC))
    w = glorot_policy()
    rs = np.random.RandomState(1)
    w[17 * 512:17 * 512 + 512] = rs.randn(512).astype(np.float32) * 0.1
    o2 = 17 * 512 + 512 + 512 * 512
    w[o2:o2 + 512] = rs.randn(512).astype(np.float32) * 0.1
    w[-2:] = (0.05, -0.03)
    g.load_policy(w)
    x = (np.random.RandomState(m).rand(m, 17).astype(np.float32) * 2 - 1)
    yd = torch.full((m, 2), 7.0, device="cuda")
    g.policy_infer(torch.from_numpy(x).cuda(), yd)
    torch.cuda.synchronize()
    y = yd.cpu().numpy()
    assert np.abs(y - numpy_mlp_bf16(w, x)).max() <= 5e-4
    assert np.abs(y - numpy_mlp(w, x)).max() <= 3e-2
    g.close()


def test_policy_tc_action_does_not_depend_on_the_batch_position():
    """Every CTA of policy_tc_kernel walks W2 in the same order: the same observation gives the same bits wherever it sits in the batch,
    and whatever the batch size."""
    g = RoverBatch(make_config(abi.PRESET_TRAINING, 8, policy_precision=abi.POLICY_TC))
    g.load_policy(glorot_policy())
    rs = np.random.RandomState(11)
    x = rs.rand(3000, 17).astype(np.float32) * 2 - 1
    probe = x[5].copy()
    for pos in (0, 127, 128, 1500, 2999):
        x[pos] = probe
    yd = torch.zeros(3000, 2, device="cuda")
    g.policy_infer(torch.from_numpy(x).cuda(), yd)
    y = yd.cpu().numpy()
    for pos in (0, 127, 128, 1500, 2999):
        assert np.array_equal(y[pos], y[5]), pos
    y1 = torch.zeros(1, 2, device="cuda")
    g.policy_infer(torch.from_numpy(probe[None]).cuda(), y1)
    assert np.array_equal(y1.cpu().numpy()[0], y[5]) and np.abs(y[5]).max() > 1e-4
    g.close()


def test_policy_tc_rollout_tracks_fp32_rollout():
    """rover_training_1 batch with the tensor-core actor inside the step loop (parked list -> policy_tc_kernel ->
    policy_act in the record): contact counts and the actions applied stay close to the fp32-actor batch over the
    first policy ticks (the trajectories are chaotic afterwards; statistics are compared, not states)."""
    n = 256
    mk = lambda prec: make_config(abi.PRESET_TRAINING, n, policy_precision=prec, exploration=0)  # noqa: E731
    ga, gb = RoverBatch(mk(abi.POLICY_TC)), RoverBatch(mk(abi.POLICY_FP32))
    w = glorot_policy()
    ga.load_policy(w); gb.load_policy(w)
    ga.step(1450); gb.step(1450)   # IC starts at 1..1.25 s, period 0.5 s: every env has taken exactly one policy action
    sa = [ga.get_state(e) for e in range(0, n, 17)]
    sb = [gb.get_state(e) for e in range(0, n, 17)]
    da = np.array([[s.steering_rate, s.boggie_torque] for s in sa])
    db = np.array([[s.steering_rate, s.boggie_torque] for s in sb])
    assert np.all(np.isfinite(da))
    assert np.abs(da[:, 0] - db[:, 0]).max() <= 15 * 3e-2 and np.abs(da[:, 1] - db[:, 1]).max() <= 25 * 3e-2
    assert np.abs(da).max() > 1e-3  # the actor really drove the setpoints
    ga.close(); gb.close()


# ----------------------------------------------------------------------------------------------------------------------
# 4. experience tuples and episode rules (src/rover_1_tf.cc:111-199, src/rover_training_1.cc:212-224)
# ----------------------------------------------------------------------------------------------------------------------
def test_transitions_match_oracle_rows():
    n = 32
    g, o = make_pair("training_fast", n, transition_capacity=8 * n * 10)
    g.step(5000); o.step(5000)
    buf = torch.zeros(8 * n * 10, abi.TRANSITION_WIDTH, device="cuda")
    with pytest.raises(Exception, match="rows pending"):
        g.drain_transitions(buf, 3)               # too small: nothing is consumed, the error names the pending count
    cnt = g.drain_transitions(buf, buf.shape[0])
    assert g.last_dropped == 0 and g.drain_transitions(buf, buf.shape[0]) == 0
    rows_o = o.drain_transitions()
    assert cnt == rows_o.shape[0] and cnt > n
    rows_g = buf[:cnt].cpu().numpy()
    # the GPU appends in completion order; the oracle env-major: compare as per-env ordered lists
    for e in range(n):
        a = rows_g[rows_g[:, 0] == e]; b = rows_o[rows_o[:, 0] == e]
        assert np.array_equal(a, b), "env %d" % e
    assert rows_g[:, 21].sum() > 0  # episodes ended: terminal flag patched into the last tuple (rover_training_1.cc:212-224)
    g.close(); o.close()


def test_done_flags_rewards_and_observations_match_oracle():
    n = 16
    g, o = make_pair("training_fast", n)
    obs = torch.zeros(n, 17, device="cuda"); rew = torch.zeros(n, device="cuda")
    done = torch.zeros(n, dtype=torch.uint8, device="cuda"); nc = torch.zeros(n, dtype=torch.int32, device="cuda")
    seen_done = 0
    for k in range(12):
        g.step(500); o.step(500)
        g.observe(obs, rew, done, nc)
        oo, ro, do, no = o.observe()
        assert np.array_equal(done.cpu().numpy(), do) and np.array_equal(nc.cpu().numpy(), no)
        assert np.array_equal(rew.cpu().numpy(), ro)
        assert np.allclose(obs.cpu().numpy(), oo, rtol=0, atol=1e-4)  # asin/atan2 in degrees: shared polynomials, float vs double print path
        seen_done += int(do.sum())
    assert seen_done > 0
    g.close(); o.close()


# ----------------------------------------------------------------------------------------------------------------------
# 5. size-independent properties at BASELINE sizes
# ----------------------------------------------------------------------------------------------------------------------
def test_config2_4096_envs_determinism_sharding_and_invariants():
    n, steps = 4096, 1500
    mk = lambda nn, off: SCENES["hf_groove_random"](nn, env_id_offset=off)  # noqa: E731
    a = RoverBatch(mk(n, 0)); a.step(steps)
    out_a = torch.zeros(n, 12, 13, device="cuda"); a.get_bodies(out_a)
    b = RoverBatch(mk(n, 0)); b.step(700); b.step(steps - 700)  # same trajectory regardless of how calls are split
    out_b = torch.zeros(n, 12, 13, device="cuda"); b.get_bodies(out_b)
    assert torch.equal(out_a, out_b)
    # sharding: envs [1024, 2048) stepped alone (as rank 1 of 4 would) are bit-identical to their slice of the full batch
    c = RoverBatch(mk(1024, 1024)); c.step(steps)
    out_c = torch.zeros(1024, 12, 13, device="cuda"); c.get_bodies(out_c)
    assert torch.equal(out_a[1024:2048], out_c)
    x = out_a[:, :10].cpu().numpy()
    assert np.isfinite(x).all()
    assert np.abs(np.linalg.norm(x[:, :, 3:7], axis=2) - 1).max() < 1e-5  # unit quaternions
    assert len(np.unique(x[:, 0, 1])) > n // 2  # random actions: envs diverged
    nc = torch.zeros(n, dtype=torch.int32, device="cuda"); a.observe(ncontacts=nc)
    assert int(nc.min()) >= 0 and int(nc.max()) <= a.cfg.max_contacts
    over, mx = a.contact_stats()
    assert over == 0 and mx <= a.cfg.max_contacts   # no contact was dropped anywhere in the 4096 x 1500 env-steps
    # spot-check 8 envs of the full batch against the oracle (bit-exact)
    o = oracle_lib.Oracle(mk(n, 0), "float", threads=16); o.step(steps)
    for e in (0, 1, 777, 1024, 2047, 3000, 4000, 4095):
        assert_states_equal(a.get_state(e), o.get_state(e), 10, True, "env %d" % e)
    for h in (a, b, c, o):
        h.close()


def test_rollout_host_round_trip_equals_device_calls():
    n = 512
    g1 = RoverBatch(SCENES["plane_ext"](n)); g2 = RoverBatch(SCENES["plane_ext"](n))
    act = (np.random.RandomState(0).rand(n, 2).astype(np.float32) * 2 - 1) * np.array([15, 25], np.float32)
    obs_h = np.zeros((n, 17), np.float32); rew_h = np.zeros(n, np.float32); done_h = np.zeros(n, np.uint8)
    g1.rollout_host(act, 300, obs_h, rew_h, done_h)
    g2.set_actions(torch.from_numpy(act).cuda()); g2.step(300)
    obs = torch.zeros(n, 17, device="cuda"); g2.observe(obs)
    assert np.array_equal(obs.cpu().numpy(), obs_h)
    assert g1.launch_count > 0
    g1.close(); g2.close()


def test_reset_subset_and_state_round_trip():
    n = 8
    g, o = make_pair("steps_random", n)
    g.step(1200); o.step(1200)
    g.reset([2, 5]); o.reset([2, 5])
    g.step(300); o.step(300)
    for e in range(n):
        assert_states_equal(g.get_state(e), o.get_state(e), 12, True, "env %d" % e)
    s = g.get_state(3)
    g.set_state(3, s)
    assert_states_equal(g.get_state(3), s, 12, True, "round trip")
    g.close(); o.close()
