"""The TD3 trainer half of SURVEY 8(f) n2 (tf_ode_b200/td3.py, mirror of scripts/MachineLearning/TD3.py and the loop of
scripts/rover_training_1.py:101-125).  CPU tests run the trainer on synthetic experience rows; the GPU test closes the loop with the
engine (trial -> device-resident experience -> train -> new actor blob)."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib
from tf_ode_b200 import abi, make_config

torch = pytest.importorskip("torch")
from tf_ode_b200 import td3 as T  # noqa: E402


def _rows(m, seed=0):
    rs = np.random.RandomState(seed)
    r = np.zeros((m, abi.TRANSITION_WIDTH), np.float32)
    r[:, 0] = rs.randint(0, 64, m)
    r[:, 1:18] = rs.randn(m, 17) * np.array(T.STATE_SCALE) * 0.3
    r[:, 18:20] = rs.uniform(-1, 1, (m, 2)) * np.array(T.ACTION_SCALE)
    r[:, 20] = r[:, 18] / 15.0 - 0.5 * np.abs(r[:, 19]) / 25.0   # a learnable reward
    r[:, 21] = rs.rand(m) < 0.05
    r[:, 22:39] = r[:, 1:18] + rs.randn(m, 17)
    return torch.from_numpy(r)


def test_actor_blob_is_the_engine_layout():
    """Keras [in,out] blob of the torch actor -> the oracle's MLP (the engine's fp32 contract) gives the torch actor's own outputs."""
    torch.manual_seed(3)
    t = T.TD3(device="cpu", seed=3)
    blob = T.actor_blob(t.actor)
    o = oracle_lib.Oracle(make_config(abi.PRESET_TRAINING, 1), "float")
    o.load_policy(blob)
    x = (np.random.RandomState(1).rand(50, 17).astype(np.float32) * 2 - 1)
    y = np.zeros((50, 2), np.float32)
    o.L.orc_mlp_infer(o.h, x.ctypes.data_as(C.c_void_p), 50, y.ctypes.data_as(C.c_void_p))
    with torch.no_grad():
        yt = t.actor(torch.from_numpy(x)).numpy()
    assert np.abs(y - yt).max() < 2e-5 and np.abs(yt).max() > 1e-3
    o.close()


def test_replay_ring_wraps_like_a_bounded_deque():
    ring = T.ReplayRing(100, torch.device("cpu"))
    a, b = _rows(70, 1), _rows(70, 2)
    ring.extend_rows(a); ring.extend_rows(b)
    assert len(ring) == 100 and ring.head == 40
    assert torch.equal(ring.s[:40], b[30:, 1:18]) and torch.equal(ring.s[70:], b[:30, 1:18]) and torch.equal(ring.s[40:70], a[40:, 1:18])
    ring.extend_rows(_rows(250, 3))     # more than the capacity at once: the newest `cap` rows survive
    assert len(ring) == 100 and torch.equal(ring.r[ring.head - 1 if ring.head else 99], _rows(250, 3)[-1, 20:21])


def test_td3_update_rules():
    t = T.TD3(device="cpu", seed=5, minibatch_size=64)
    assert t.train(3) == 0.0 and t.n_iter == 0              # TD3.py:313-314: fewer samples than a minibatch -> nothing happens
    t.replay_buffer.extend_rows(_rows(4000, 7))
    a0 = [p.clone() for p in t.actor.parameters()]
    at0 = [p.clone() for p in t.actor_target.parameters()]
    lq_first = t.train(20)
    assert t.n_iter == 20 and np.isfinite(lq_first) and lq_first > 0
    # soft target update: target = tau * new + (1 - tau) * old stays close to the old target, the live actor has moved more
    d_live = max(float((p - q).abs().max()) for p, q in zip(t.actor.parameters(), a0))
    d_tgt = max(float((p - q).abs().max()) for p, q in zip(t.actor_target.parameters(), at0))
    assert 0 < d_tgt < d_live
    lq_later = t.train(150)
    assert lq_later < lq_first                               # the Bellman residual on a learnable reward goes down
    act = t.get_action(np.zeros(17, np.float32))
    assert act.shape == (2,) and abs(act[0]) <= 15 and abs(act[1]) <= 25


@pytest.mark.gpu
def test_training_loop_closes_on_the_gpu():
    """trial on the engine -> experience stays on the device -> TD3 iterations -> new actor blob into the engine, twice."""
    lines = []
    t = T.TD3(device="cuda", seed=11)
    hist = T.training_loop(t, n_envs=64, max_episodes=128, iter_per_ep=1, log=lines.append, speedf=0.5, x_goal=0.6)
    assert len(hist) == 2 and hist[0]["rows"] > 64 and hist[1]["buffer"] == hist[0]["rows"] + hist[1]["rows"]
    assert hist[1]["it"] == 128 and np.isfinite(hist[1]["LQ"]) and hist[1]["LQ"] > 0
    assert lines[0].startswith("It 64 | Ep 64 | Bs ")        # the reference's progress line (rover_training_1.py:123-124)
    s = t.replay_buffer.s[: len(t.replay_buffer)]
    assert float(s.abs().max()) > 1 and bool(torch.isfinite(s).all())
    assert float(t.replay_buffer.t.sum()) >= 64               # every episode's last tuple is terminal (rover_training_1.cc:212-224)
