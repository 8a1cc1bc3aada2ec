"""Parity at the sizes bench.py times (BASELINE.json configs 2-5) and the long-horizon contract of SURVEY 8(c).

* every benchmarked configuration at its benchmark size: sampled environments of the full batch are re-run one by one on
  oracle<float> (same global id, same calls) and must be bit-equal; the contact-capacity audit must read zero;
* episode statistics over >= 4096 COMPLETE episodes, GPU fp32 against oracle<double> (the reference-equivalent arithmetic; its 4096
  episodes per scene are the committed fixture tests/golden/long_horizon_*.npz, pinned to the live oracle by tests/test_oracle_kat.py):
  episode return mean and sigma, success rate, episode length, each within 2 standard errors.
Run with `-m gpu` on a B200; everything goes through the C ABI.
"""
import os
import sys

import numpy as np
import pytest

import bench
import oracle_lib
from tf_ode_b200 import RoverBatch, abi, glorot_policy, make_config, reference_heightmap

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _oracle_env(workload, gid, prec="float", **over):
    cfg = bench.make_cfg(workload, 1, env_id_offset=gid, **over)
    o = oracle_lib.Oracle(cfg, prec, threads=1)
    if cfg.controller == abi.CTRL_MLP:
        o.load_policy(glorot_policy())
    return o


def _assert_env_equal(g, e, o, nb, tag):
    sg, so = g.get_state(e), o.get_state(0)
    bg, bo = oracle_lib.state_bodies(sg)[:nb], oracle_lib.state_bodies(so)[:nb]
    for f in ("lcg_seed", "n_contacts", "tick", "time_steps", "episode", "episodes_done", "explore", "ic_activated"):
        assert getattr(sg, f) == getattr(so, f), "%s: %s %r != %r" % (tag, f, getattr(sg, f), getattr(so, f))
    assert np.array_equal(bg, bo), "%s: bodies differ, max |d| = %.3e" % (tag, np.abs(bg - bo).max())
    assert sg.total_reward == so.total_reward and sg.steering_rate == so.steering_rate and sg.boggie_torque == so.boggie_torque, tag


# ----------------------------------------------------------------------------------------------------------------------
# 1. the benchmarked configurations at their benchmark sizes
# ----------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("workload,n,steps", [("config2", 4096, 2500), ("config3", 16384, 2000), ("config4", 8192, 2000), ("config5", 131072, 1200)])
def test_benchmarked_config_at_full_size_matches_oracle_bit_exact(workload, n, steps):
    """config 3 runs the 17-512-512-2 actor fused inside step_kernel (2-3 policy ticks per environment in 2000 sub-steps, e-greedy on);
    config 4 carries the per-environment amplitude / rate jitter; config 5 is the 400 x 400 map."""
    off = 3 * n   # a shard in the middle of a larger job: global ids off .. off + n
    g = RoverBatch(bench.make_cfg(workload, n, env_id_offset=off))
    if g.cfg.controller == abi.CTRL_MLP:
        g.load_policy(glorot_policy())
    nb = 12 if g.cfg.scene == abi.SCENE_STEP_BOXES else 10
    act = bench.action_tape(0, 0, n, off)
    g.set_actions(torch.from_numpy(act).cuda())
    g.step(steps // 2); g.step(steps - steps // 2)
    ids = sorted({0, 1, n // 2, n - 1} | set(int(x) for x in np.random.RandomState(n).randint(0, n, size=8)))
    for e in ids:
        o = _oracle_env(workload, off + e)
        o.set_actions(act[e:e + 1])
        o.step(steps)
        _assert_env_equal(g, e, o, nb, "%s env %d" % (workload, off + e))
        o.close()
    over, mx = g.contact_stats()
    assert over == 0 and 0 < mx <= g.cfg.max_contacts, "contacts were dropped: %d sub-steps over capacity, up to %d generated" % (over, mx)
    out = torch.zeros(n, 12, 13, device="cuda"); g.get_bodies(out)
    x = out[:, :10].cpu().numpy()
    assert np.isfinite(x).all() and np.abs(np.linalg.norm(x[:, :, 3:7], axis=2) - 1).max() < 1e-5
    g.close()


def test_contact_capacity_audit_counts_dropped_contacts():
    """With the capacity forced below what the scene generates the audit must say so (and the oracle applies the same cut)."""
    n = 64
    mk = lambda: make_config(abi.PRESET_MT_RANDOM, n, max_contacts=6, heightfield=(reference_heightmap("rock_groove"), 3.0, 3.0, (1.0, 0.0, -0.3)))  # noqa: E731
    g = RoverBatch(mk()); o = oracle_lib.Oracle(mk(), "float", threads=8)
    g.step(800); o.step(800)
    over, mx = g.contact_stats()
    h = o.contact_hist()
    assert over == int(h[7:].sum()) and over > 0 and mx == int(np.nonzero(h)[0].max()) and mx > 6
    for e in (0, 31, 63):
        assert np.array_equal(oracle_lib.state_bodies(g.get_state(e)), oracle_lib.state_bodies(o.get_state(e)))
    over2, _ = g.contact_stats(reset=True)
    assert over2 == over and g.contact_stats() == (0, 0)
    g.close(); o.close()


def test_fused_actor_equals_parked_fp32_actor_and_tc_actor_stays_close():
    """The actor fused into step_kernel (TFODE_POLICY_FP32) against the oracle's own fp32 MLP over many policy ticks, and the parked
    tensor-core path (TFODE_POLICY_TC) against it on the first tick's actions (bf16 operands: 3e-2 of full scale)."""
    n = 256
    mk = lambda prec: make_config(abi.PRESET_TRAINING, n, policy_precision=prec, exploration=0)  # noqa: E731
    gf, gt = RoverBatch(mk(abi.POLICY_FP32)), RoverBatch(mk(abi.POLICY_TC))
    o = oracle_lib.Oracle(mk(abi.POLICY_FP32), "float", threads=8)
    w = glorot_policy()
    for h in (gf, gt, o):
        h.load_policy(w)
    l0 = gf.launch_count
    gf.step(1200); gt.step(1200); o.step(1200)   # IC starts at 0.75..1.25 s: up to one policy action each
    assert gf.launch_count - l0 == 1              # fused: ONE launch per call, no parking
    da = np.array([[gf.get_state(e).steering_rate, gf.get_state(e).boggie_torque] for e in range(0, n, 5)])
    db = np.array([[gt.get_state(e).steering_rate, gt.get_state(e).boggie_torque] for e in range(0, n, 5)])
    do = np.array([[o.get_state(e).steering_rate, o.get_state(e).boggie_torque] for e in range(0, n, 5)])
    assert np.array_equal(da, do) and np.abs(da).max() > 1e-3
    assert np.abs(da[:, 0] - db[:, 0]).max() <= 15 * 3e-2 and np.abs(da[:, 1] - db[:, 1]).max() <= 25 * 3e-2
    gf.step(3800); o.step(3800)                   # ~8 more ticks per environment
    for e in range(0, n, 37):
        assert np.array_equal(oracle_lib.state_bodies(gf.get_state(e)), oracle_lib.state_bodies(o.get_state(e))), e
        assert gf.get_state(e).tick == o.get_state(e).tick >= 7
    for h in (gf, gt, o):
        h.close()


# ----------------------------------------------------------------------------------------------------------------------
# 2. long horizon: statistics over complete episodes, GPU fp32 vs oracle<double>
# ----------------------------------------------------------------------------------------------------------------------
def _episodes(h, n):
    s = [h.get_state(e) for e in range(n)]
    return (np.array([x.total_reward for x in s]), np.array([float(x.time_steps) for x in s]), np.array([x.term_reason for x in s]),
            np.array([x.episodes_done for x in s]), np.array([[x.pos[0][0], x.pos[0][1]] for x in s]))


def _se_mean(a, b):
    return np.sqrt(a.var(ddof=1) / a.size + b.var(ddof=1) / b.size)


def _se_sigma(a):
    m = a.mean(); s2 = a.var()
    m4 = ((a - m) ** 4).mean()
    return np.sqrt(max(m4 - s2 * s2, 1e-300) / (4 * max(s2, 1e-300) * a.size))   # delta method, no normality assumed


def _compare(name, a, b, report):
    dm, sem = abs(a.mean() - b.mean()), _se_mean(a, b) + 1e-12
    ds, ses = abs(a.std() - b.std()), np.hypot(_se_sigma(a), _se_sigma(b)) + 1e-12
    report.append("%-14s GPU mean %.6g sigma %.6g | oracle<double> mean %.6g sigma %.6g | d mean %.2f SE, d sigma %.2f SE" % (name, a.mean(), a.std(), b.mean(), b.std(), dm / sem, ds / ses))
    assert dm <= 2 * sem, report[-1]
    assert ds <= 2 * ses, report[-1]


def _golden_episodes(scene):
    d = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "long_horizon_%s.npz" % scene))
    return d["total_reward"], d["time_steps"].astype(float), d["term_reason"], d["episodes_done"], d["pos_xy"]


@pytest.mark.parametrize("scene", ["training_fast", "config2"])
def test_long_horizon_episode_statistics_against_double_oracle(scene):
    """>= 4096 complete episodes per arm (auto_reset off: one episode per environment, run to the 60 s timeout at the latest).
    GPU fp32 and oracle<double> share the per-episode random draws (yaw, IC start, exploration, actions: Philox keyed on the global
    id), so the comparison is paired in its inputs and chaotic only through the arithmetic.  The criterion is SURVEY 8(c)'s: episode
    return mean and sigma (MLP scene), success rate, episode length, each within 2 standard errors.
    The oracle<double> arm is the committed fixture tests/golden/long_horizon_<scene>.npz (tools/make_long_horizon_golden.py: the same
    configuration, 4096 episodes; tests/test_oracle_kat.py re-runs its first rows on the CPU) -- three minutes of host time per scene
    that the GPU box need not repeat.  TFODE_LONG_HORIZON_LIVE=1 runs the oracle live instead and requires it equal to the fixture."""
    sys.path.insert(0, os.path.join(bench.ROOT, "tools"))
    import make_long_horizon_golden as mk
    n = mk.N_EPISODES
    g = RoverBatch(mk.scene_config(scene, n))
    if g.cfg.controller == abi.CTRL_MLP:
        g.load_policy(glorot_policy())
    for _ in range(mk.SUBSTEPS // 1000):
        g.step(1000)
    rg, lg, tg, eg, pg = _episodes(g, n)
    ro, lo, to, eo, po = _golden_episodes(scene)
    if os.environ.get("TFODE_LONG_HORIZON_LIVE") == "1":
        live = mk.oracle_episodes(scene, n, os.cpu_count() or 8)
        assert np.array_equal(live["total_reward"], ro) and np.array_equal(live["time_steps"].astype(float), lo) and np.array_equal(live["term_reason"], to)
    assert ro.size == n and (eg == 1).all() and (eo == 1).all(), "every environment must have finished exactly one episode"
    report = []
    if scene == "training_fast":
        _compare("episode return", rg, ro, report)
    _compare("episode length", lg, lo, report)
    _compare("final x", pg[:, 0], po[:, 0], report)
    sg, so = (tg == abi.TERM_X_GOAL).astype(float), (to == abi.TERM_X_GOAL).astype(float)
    se = np.sqrt(sg.mean() * (1 - sg.mean()) / n + so.mean() * (1 - so.mean()) / n) + 1e-12
    report.append("%-14s GPU %.4f | oracle<double> %.4f | d %.2f SE" % ("success rate", sg.mean(), so.mean(), abs(sg.mean() - so.mean()) / se))
    assert abs(sg.mean() - so.mean()) <= 2 * se, report[-1]
    for reason in (abi.TERM_TIMEOUT, abi.TERM_Y_MAX, abi.TERM_UPSIDE_DOWN):
        a, b = (tg == reason).mean(), (to == reason).mean()
        report.append("reason %d       GPU %.4f | oracle<double> %.4f" % (reason, a, b))
    os.makedirs(os.path.join(bench.ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(bench.ROOT, "gpurun_out", "long_horizon_%s.txt" % scene), "w") as f:
        f.write("\n".join(report) + "\n")
    over, mx = g.contact_stats()
    assert over == 0, "contacts dropped in %d sub-steps (max generated %d)" % (over, mx)
    g.close()


def test_external_commands_on_crawlers_clamp_at_each_rover_own_amplitude():
    """SetBoggieTorque clamps at boggie_max_torque, which Crawler_1's constructor sets to ITS torque amplitude (src/rover_1_crawlers.cc:22) --
    jittered per environment in config 4.  Commands re-issued every 100 sub-steps, as bench.py does (found by bench.py --verify)."""
    n = 256
    g = RoverBatch(bench.make_cfg("config4", n)); o = oracle_lib.Oracle(bench.make_cfg("config4", n), "float", threads=8)
    amps = np.array([g.get_state(e).crawler_amp for e in range(n)])
    assert amps.min() < 21 and amps.max() > 29
    for k in range(10):
        a = bench.action_tape(0, k, n)
        g.set_actions(torch.from_numpy(a).cuda()); o.set_actions(a)
        if k == 0:
            tq = np.array([g.get_state(e).boggie_torque for e in range(n)])
            assert np.array_equal(tq, np.array([o.get_state(e).boggie_torque for e in range(n)])) and (np.abs(tq) <= amps + 1e-9).all()
            assert (np.abs(a[:, 1]) > amps).any()    # the tape does ask for more than some rovers allow
        g.step(100); o.step(100)
    for e in range(n):
        assert np.array_equal(oracle_lib.state_bodies(g.get_state(e))[:10], oracle_lib.state_bodies(o.get_state(e))[:10]), e
    g.close(); o.close()


def test_persistent_launch_equals_plain_launch(monkeypatch):
    """Batches larger than one wave of resident CTAs are stepped by a persistent grid pulling (group, 25-sub-step slice) items; the records it
    leaves are bit-identical to the plain one-CTA-per-group launch (TFODE_PERSISTENT=0), whatever the call length and the batch size."""
    for workload, n, calls in (("config2", 4096, (100, 37, 263)), ("config4", 8192 + 24, (500,)), ("config3", 3000, (130, 270))):
        outs = []
        for persistent in ("1", "0"):
            monkeypatch.setenv("TFODE_PERSISTENT", persistent)
            g = RoverBatch(bench.make_cfg(workload, n))
            if g.cfg.controller == abi.CTRL_MLP:
                g.load_policy(glorot_policy())
            g.set_actions(torch.from_numpy(bench.action_tape(0, 1, n)).cuda())
            for k in calls:
                g.step(k)
            out = torch.zeros(n, 12, 13, device="cuda"); g.get_bodies(out)
            rows = torch.zeros(n, abi.ROLLOUT_WIDTH, device="cuda"); g.pack_rollout(rows)
            outs.append((out.clone(), rows.clone(), g.get_state(n - 1).lcg_seed, g.get_state(5).time_steps))
            g.close()
        assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1]) and outs[0][2:] == outs[1][2:], workload
        assert outs[0][3] == sum(calls)
