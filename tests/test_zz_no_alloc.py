"""tfode_batch_step_dantzig_static (include/tfode_batch.h): the stepping call that allocates nothing -- the dense solver's LCP work
arrays live in a file-scope `static __device__` array of the library instead of cudaMalloc'd memory.  No counterpart in the reference,
so the checker is the existing entry points: set_actions + tfode_batch_step +
get_bodies on a TFODE_SOLVER_DANTZIG handle.  Both run the same kernel on the same inputs, so the tolerance is the one the Dantzig
parity tests use: bit-exact.  Collected last on purpose (pytest -x must not hide other tests behind this one); passes
alone in a fresh process and after the rest of the suite.
"""
import numpy as np
import pytest

from tf_ode_b200 import RoverBatch, abi, make_config, reference_heightmap

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _cfg(n, solver, **over):
    return make_config(abi.PRESET_MT_RANDOM, n, solver=solver, env_id_offset=1000,
                       heightfield=(reference_heightmap("rock_groove"), 3.0, 3.0, (1.0, 0.0, -0.3)), **over)


def test_static_scratch_step_matches_allocating_step_on_two_calls_in_a_row():
    n = abi.NOALLOC_MAX_ENVS                      # the stated maximum: the whole static array is in use
    ref = RoverBatch(_cfg(n, abi.SOLVER_DANTZIG))    # existing path: scratch cudaMalloc'd at creation
    sta = RoverBatch(_cfg(n, abi.SOLVER_QUICKSTEP))  # no dense-solver scratch in this handle at all
    rs = np.random.RandomState(5)
    tape = [torch.from_numpy(np.stack([rs.uniform(-15, 15, n), rs.uniform(-25, 25, n)], 1).astype(np.float32)).cuda() for _ in range(2)]
    out_ref = torch.zeros(n, 12, 13, device="cuda"); out_sta = torch.full((n, 12, 13), float("nan"), device="cuda")
    for call, steps in enumerate((400, 300)):     # two calls in a row: the second finds the first one's contents in the scratch
        act = tape[call]
        ref.set_actions(act); ref.step(steps); ref.get_bodies(out_ref)
        l0 = sta.launch_count
        free0 = torch.cuda.mem_get_info()[0]
        sta.step_dantzig_static(act, steps, out_sta)
        torch.cuda.synchronize()
        assert sta.launch_count - l0 == 3
        if call == 1:                             # the first call may load the module that holds the array; the second moves nothing
            assert torch.cuda.mem_get_info()[0] == free0
        a, b = out_ref.cpu().numpy(), out_sta.cpu().numpy()
        assert np.isfinite(a).all() and np.array_equal(a, b), "call %d: max |d| = %.3e" % (call, np.nanmax(np.abs(a - b)))
        for e in (0, n // 2, n - 1):
            sr, ss = ref.get_state(e), sta.get_state(e)
            assert (sr.lcg_seed, sr.n_contacts, sr.time_steps) == (ss.lcg_seed, ss.n_contacts, ss.time_steps)
    assert np.abs(out_sta[:, 0, 0].cpu().numpy() - out_sta[0, 0, 0].item()).max() > 1e-4   # the rovers did move apart
    assert ref.lcp_failures() == sta.lcp_failures()
    # the Dantzig steps differ from QuickStep ones: the static path really ran the dense solver
    qs = RoverBatch(_cfg(n, abi.SOLVER_QUICKSTEP))
    qs.set_actions(tape[0]); qs.step(400); qs.set_actions(tape[1]); qs.step(300)
    out_qs = torch.zeros(n, 12, 13, device="cuda"); qs.get_bodies(out_qs)
    assert not np.array_equal(out_qs.cpu().numpy(), out_sta.cpu().numpy())
    # NULL input / output, zero sub-steps: a no-op that still succeeds
    sta.step_dantzig_static(None, 0, None)
    for h in (ref, sta, qs):
        h.close()


def test_static_scratch_step_rejects_sizes_beyond_the_maximum():
    g = RoverBatch(make_config(abi.PRESET_SCENE_1, abi.NOALLOC_MAX_ENVS + 1))
    with pytest.raises(Exception) as ei:
        g.step_dantzig_static(None, 10, None)
    assert "static scratch holds %d" % abi.NOALLOC_MAX_ENVS in str(ei.value)
    before = torch.zeros(abi.NOALLOC_MAX_ENVS + 1, 12, 13, device="cuda"); after = torch.zeros_like(before)
    g.get_bodies(before)
    with pytest.raises(Exception):
        g.step_dantzig_static(None, 10, after)
    g.get_bodies(after)
    assert torch.equal(before, after)             # nothing ran
    g.close()
