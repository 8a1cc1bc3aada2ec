"""The dense mixed-LCP solver of the dWorldStep path (include/tfode_lcp.h) on the CPU: known answers on 1-3 row hand
cases, agreement with a brute-force enumeration of the active sets on random small problems (the solution of a P-matrix LCP is
unique), friction rows coupled to their normal through findex, and the oracle stepping a rover with TFODE_SOLVER_DANTZIG."""
import ctypes as C
import itertools

import numpy as np
import pytest

import oracle_lib
from tf_ode_b200 import abi, make_config

INF = float("inf")


def solve(A, b, lo, hi, findex=None):
    A = np.ascontiguousarray(A, np.float64); n = A.shape[0]
    b = np.ascontiguousarray(b, np.float64); lo = np.ascontiguousarray(lo, np.float64); hi = np.ascontiguousarray(hi, np.float64)
    fi = np.ascontiguousarray(findex if findex is not None else -np.ones(n), np.int32)
    x = np.zeros(n)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    st = oracle_lib.lib().orc_lcp_solve(n, p(A), p(b), p(lo), p(hi), p(fi), p(x))
    return x, st


def brute(A, b, lo, hi):
    """Enumerate every assignment (free / at lo / at hi) and return the one that satisfies the LCP conditions."""
    n = len(b)
    for states in itertools.product((0, 1, 2), repeat=n):
        if any((s == 1 and lo[i] == -INF) or (s == 2 and hi[i] == INF) for i, s in enumerate(states)):
            continue
        x = np.array([lo[i] if s == 1 else (hi[i] if s == 2 else 0.0) for i, s in enumerate(states)])
        F = [i for i, s in enumerate(states) if s == 0]
        if F:
            Bn = [i for i in range(n) if i not in F]
            rhs = b[F] - (A[np.ix_(F, Bn)] @ x[Bn] if Bn else 0)
            x[F] = np.linalg.solve(A[np.ix_(F, F)], rhs)
        w = A @ x - b
        ok = True
        for i, s in enumerate(states):
            if s == 0 and not (lo[i] - 1e-12 <= x[i] <= hi[i] + 1e-12 and abs(w[i]) < 1e-9):
                ok = False
            if s == 1 and w[i] < -1e-10:
                ok = False
            if s == 2 and w[i] > 1e-10:
                ok = False
        if ok:
            return x
    return None


def test_one_row_known_answers():
    assert solve([[2.0]], [1.0], [0.0], [INF])[0][0] == pytest.approx(0.5)          # pulled into contact
    assert solve([[2.0]], [-1.0], [0.0], [INF])[0][0] == 0.0                        # separating: lambda = 0, w = 1
    assert solve([[2.0]], [10.0], [-3.0], [3.0])[0][0] == 3.0                       # motor saturates at fmax
    assert solve([[4.0]], [-2.0], [-INF], [INF])[0][0] == pytest.approx(-0.5)       # equality row


def test_two_and_three_rows_against_hand_solutions():
    A = np.array([[2.0, 1.0], [1.0, 2.0]])
    x, st = solve(A, [3.0, 3.0], [0, 0], [INF, INF]); assert st == 0 and np.allclose(x, [1, 1])
    x, st = solve(A, [3.0, -3.0], [0, 0], [INF, INF]); assert st == 0 and np.allclose(x, [1.5, 0])     # second row lifts off
    x, st = solve(A, [3.0, 3.0], [0, 0], [0.5, INF]); assert st == 0 and np.allclose(x, [0.5, 1.25])   # first row clamps at hi
    A3 = np.array([[4.0, 1, 0], [1, 3, 1], [0, 1, 2]])
    x, st = solve(A3, [1.0, 2.0, -5.0], [-INF, 0, 0], [INF, INF, INF])
    assert st == 0 and np.allclose(x, brute(A3, np.array([1.0, 2.0, -5.0]), [-INF, 0, 0], [INF, INF, INF]))


def test_random_small_problems_match_brute_force():
    rs = np.random.RandomState(12)
    for trial in range(300):
        n = rs.randint(2, 7)
        M = rs.randn(n, n + 2)
        A = M @ M.T + 0.05 * np.eye(n)
        b = rs.randn(n) * 2
        kind = rs.randint(0, 4, n)
        lo = np.where(kind == 0, -INF, np.where(kind == 1, 0.0, -rs.rand(n) * 2))
        hi = np.where(kind == 0, INF, np.where(kind == 1, INF, rs.rand(n) * 2))
        x, st = solve(A, b, lo, hi)
        ref = brute(A, b, lo, hi)
        assert st == 0 and ref is not None
        assert np.allclose(x, ref, atol=1e-8), (trial, x, ref)


def test_friction_rows_follow_their_normal_row():
    """Contact triple (normal, friction 1, friction 2) with dContactApprox1 bounds: |lambda_t| <= mu * lambda_n."""
    A = np.diag([2.0, 1.0, 1.0]) + 0.1
    mu = 0.5
    x, st = solve(A, [4.0, 5.0, -0.2], [0, -mu, -mu], [INF, mu, mu], findex=[-1, 0, 0])
    assert st == 0
    # the normal row is solved with the friction rows at zero first: lambda_n = 4 / 2.1, bounds +-mu * lambda_n; later pivots may move
    # lambda_n, the bounds stay (lcp.cpp sets them once, when the first friction row is reached)
    bound = mu * 4.0 / 2.1
    assert x[1] == pytest.approx(bound) and abs(x[2]) < bound
    w = A @ x - np.array([4.0, 5.0, -0.2])
    assert abs(w[0]) < 1e-9 and w[1] < 0 and abs(w[2]) < 1e-9
    x, st = solve(A, [-4.0, 5.0, -0.2], [0, -mu, -mu], [INF, mu, mu], findex=[-1, 0, 0])
    assert st == 0 and np.array_equal(x, [0, 0, 0])          # no normal force, no friction


def test_oracle_steps_a_rover_with_the_dantzig_solver():
    """dWorldStep path on the oracle: no stalled drive, the rover settles and cruises like under QuickStep (the two solvers agree on
    a well-conditioned scene to the accuracy of 20 SOR sweeps)."""
    mk = lambda s: make_config(abi.PRESET_SCENE_1, 1, scene=abi.SCENE_PLANE, solver=s)  # noqa: E731
    od, oq = oracle_lib.Oracle(mk(abi.SOLVER_DANTZIG), "double"), oracle_lib.Oracle(mk(abi.SOLVER_QUICKSTEP), "double")
    od.step(1500); oq.step(1500)
    bd, bq = od.bodies(0)[:10], oq.bodies(0)[:10]
    assert oracle_lib.lib().orc_lcp_failures(od.h) == 0
    assert np.isfinite(bd).all() and abs(bd[0, 2] - bq[0, 2]) < 2e-3           # chassis height
    assert 0.02 < bd[0, 0] - 0.1095 < 0.08 and abs(bd[0, 0] - bq[0, 0]) < 5e-3  # ~0.04 m/s for ~1.25 s after the ramp
    assert od.get_state(0).lcg_seed == 0 and oq.get_state(0).lcg_seed != 0       # dWorldStep draws no random numbers
    assert od.get_state(0).n_contacts == oq.get_state(0).n_contacts == 8 or od.get_state(0).n_contacts in (4, 6, 8)
    od.close(); oq.close()
