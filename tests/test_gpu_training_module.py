"""trial()/eval() of tf_ode_b200.training_module (the batched rover_training_1_module) vs the CPU oracle.  `-m gpu`."""
import io

import numpy as np
import pytest

import oracle_lib
from tf_ode_b200 import abi, glorot_policy, make_config, training_module

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

FAST = dict(speedf=0.5, x_goal=0.6)   # episodes of a few seconds instead of a minute


def _oracle_rows(n, exploration):
    """Every episode to its end (goal, y_max, upside-down or the 60 s timeout), like training_module._run_episodes."""
    cfg = make_config(abi.PRESET_TRAINING, n, auto_reset=0, exploration=int(exploration), transition_capacity=n * 122, **FAST)
    o = oracle_lib.Oracle(cfg, "float", threads=8)
    o.load_policy(glorot_policy())
    for _ in range(61):
        o.step(1000)
        if all(o.get_state(e).episodes_done == 1 for e in range(n)):
            break
    rows = o.drain_transitions()
    states = [o.get_state(e) for e in range(n)]
    o.close()
    return rows, states


def test_trial_returns_the_reference_tuples_and_matches_the_oracle():
    n = 12
    exp = training_module.trial(glorot_policy(), n_episodes=n, **FAST)
    rows_o, states = _oracle_rows(n, True)
    assert all(s.episodes_done == 1 for s in states)
    want = training_module.rows_to_experience(rows_o, n)
    assert len(exp) == n
    for e in range(n):
        assert len(exp[e]) == len(want[e]) and len(exp[e]) >= 1, "env %d" % e
        for a, b in zip(exp[e], want[e]):
            assert a == b
        s, act, r, terminal, s2 = exp[e][-1]
        assert len(s) == 17 and len(s2) == 17 and len(act) == 2 and isinstance(r, float)
        assert terminal is True and all(t[3] is False for t in exp[e][:-1])   # only the last tuple is terminal (:212-224)
    single = training_module.trial(glorot_policy(), n_episodes=1, **FAST)
    assert single == exp[0]   # env 0 of a batch is the same episode as a batch of one (streams are keyed on the env id)


def test_eval_prints_the_reference_result_line():
    n = 4
    buf = io.StringIO()
    res = training_module.eval(glorot_policy(), n_episodes=n, colour=False, file=buf, **FAST)
    _, states = _oracle_rows(n, False)
    lines = buf.getvalue().strip().splitlines()
    assert len(lines) == n
    for e, (line, s) in enumerate(zip(lines, states)):
        assert line.startswith("[Success] t ") or line.startswith("[Failure] t ")
        assert " | x " in line and " | y " in line and " | Rmoy " in line
        ok, t, x, y, rmoy = res[e]
        assert t == pytest.approx(float(np.float32(s.time_steps) * np.float32(0.001)))
        assert rmoy == pytest.approx(s.total_reward / t, rel=1e-12)
        assert ok == (abs(x) >= 0.6)


def test_trial_takes_the_reference_argument_a_saved_model_directory(tmp_path):
    """trial( session_dir + '/actor' ) (scripts/rover_training_1.py:121): the directory gives the same experience as its weight blob."""
    from test_formats import write_bundle
    w = glorot_policy()
    t, o = {}, 0
    for l, (i, j) in enumerate(((17, 512), (512, 512), (512, 2))):
        t["layer_with_weights-%d/kernel/.ATTRIBUTES/VARIABLE_VALUE" % l] = (w[o:o + i * j].reshape(i, j).copy(), 1); o += i * j
        t["layer_with_weights-%d/bias/.ATTRIBUTES/VARIABLE_VALUE" % l] = (w[o:o + j].copy(), 1); o += j
    write_bundle(str(tmp_path / "actor" / "variables" / "variables"), t)
    a = training_module.trial(str(tmp_path / "actor"), n_episodes=1, **FAST)
    b = training_module.trial(w, n_episodes=1, **FAST)
    assert len(a) == len(b) > 0
    for ra, rb in zip(a, b):
        assert ra[0] == rb[0] and ra[1] == rb[1] and ra[2] == rb[2] and ra[3] == rb[3] and ra[4] == rb[4]
