"""GPU model-tree controller (Rover_1_mt) and the reference's own terrains vs the CPU oracle.  `-m gpu`, through the C ABI."""
import os

import numpy as np
import pytest

import oracle_lib
from tf_ode_b200 import PRESET_MT_TREE, RoverBatch, TfodeError, abi, make_config, reference_heightmap
from test_gpu_parity import assert_states_equal, nbodies

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
Y1 = open(os.path.join(GOLD, "model_tree_1.yaml")).read()
Y2 = open(os.path.join(GOLD, "model_tree_2.yaml")).read()
YP = open(os.path.join(GOLD, "model_tree_poly2.yaml")).read()
IO = np.load(os.path.join(GOLD, "model_tree_io.npz"))


def tree_pair(n=1, prec="float", trees=(Y1, Y2), degree=1, **kw):
    mk = lambda: make_config(PRESET_MT_TREE, n, **kw)  # noqa: E731
    g, o = RoverBatch(mk()), oracle_lib.Oracle(mk(), prec)
    g.load_model_trees(trees[0], trees[1], degree=degree)
    o.load_model_trees(trees[0], trees[1], degree=degree)
    return g, o


@pytest.mark.parametrize("degree,trees", [(1, (Y1, Y2)), (2, (YP, YP))])
def test_model_tree_infer_is_bit_exact_with_oracle_and_matches_reference_python(degree, trees):
    g, o = tree_pair(1, "double", trees, degree)
    x = torch.from_numpy(IO["X"]).cuda()
    act = torch.zeros(x.shape[0], 2, dtype=torch.float64, device="cuda")
    node = torch.zeros(x.shape[0], 2, dtype=torch.int32, device="cuda")
    g.model_tree_infer(x, act, node)
    a, nd = act.cpu().numpy(), node.cpu().numpy()
    for which in (0, 1):
        y, n_o = o.model_tree_predict(which, IO["X"])
        assert np.array_equal(nd[:, which], n_o)
        assert np.array_equal(a[:, which], y)      # same fp64 operations in the same order
    name = ("1", "2") if degree == 1 else ("poly2", "poly2")
    for which in (0, 1):                            # the reference's own model_tree.py predictions (tools/make_golden.py)
        assert np.array_equal(nd[:, which], IO["node_" + name[which]])
        assert np.allclose(a[:, which], IO["y_" + name[which]], rtol=1e-10, atol=1e-10)
    g.close(); o.close()


def test_model_tree_load_rejects_what_the_reference_rejects():
    g = RoverBatch(make_config(PRESET_MT_TREE, 1))
    with pytest.raises(TfodeError, match="needs tfode_batch_load_model_trees"):
        g.step(1)
    with pytest.raises(TfodeError, match="does not fit the number of parameters"):   # model_tree.cc:50-51
        g.load_model_trees(Y1, Y2, degree=2)
    with pytest.raises(TfodeError, match="instead of the 2 required"):               # model_tree.cc:33-35
        g.load_model_trees("1:\n  terminal: false\n  split params: [0, 1, 2]\n", Y2)
    with pytest.raises(TfodeError, match="degree"):
        g.load_model_trees(Y1, Y2, degree=3)
    g.close()


@pytest.mark.parametrize("kw,steps", [(dict(), 4000), (dict(step_yaw_deg=4.0, speedf=0.3), 5000)])
def test_model_tree_trajectory_is_bit_exact(kw, steps):
    """scene_1_mt as shipped (step boxes, Rover_1_mt at 2 Hz) free-running against oracle<float>."""
    g, o = tree_pair(1, "float", **kw)
    nb = nbodies(g.cfg)
    k, chunk = 0, 1
    while k < steps:
        g.step(chunk); o.step(chunk); k += chunk
        assert_states_equal(g.get_state(0), o.get_state(0), nb, True, "mt step %d" % k)
        chunk = 1 if k < 100 else 41
    assert g.get_state(0).tick >= 5
    g.close(); o.close()


def test_model_tree_batch_matches_oracle_and_reports_terminal_nodes():
    n = 64
    hm = reference_heightmap("rock_groove")
    g, o = tree_pair(n, "float", heightfield=(hm, 3.0, 3.0, (1.0, 0.0, -0.3)), speedf=0.4, ic_start_jitter=0.25)
    nc = torch.zeros(n, dtype=torch.int32, device="cuda")
    for k in range(30):
        g.step(50); o.step(50)
        g.observe(ncontacts=nc)
        assert np.array_equal(nc.cpu().numpy(), o.observe()[3]), "step %d" % ((k + 1) * 50)
    for e in range(0, n, 5):
        assert_states_equal(g.get_state(e), o.get_state(e), 10, True, "env %d" % e)
    nodes = torch.zeros(n, 2, dtype=torch.int32, device="cuda")
    g.get_tree_nodes(nodes)
    nodes = nodes.cpu().numpy()
    terminal_ids = {4, 5, 7, 8, 11, 12, 14, 15}   # depth-3 full binary tree, ids handed out depth-first
    assert set(nodes.ravel().tolist()) <= terminal_ids
    g.close(); o.close()


@pytest.mark.parametrize("kind,pos,speed", [("rock_groove", (1.0, 0.0, -0.3), 0.5), ("rock_step", (2.0, 0.0, -0.01), 0.5)])
def test_reference_terrain_trajectory_is_bit_exact(kind, pos, speed):
    """The heightfield scenes of src/scene_1.cc:60-62 / scene_1_mt.cc:65-67 on the reference's own maps."""
    hm = reference_heightmap(kind)
    mk = lambda: make_config(abi.PRESET_MT_RANDOM, 1, heightfield=(hm, 3.0, 3.0, pos), speedf=speed, x_goal=2.5)  # noqa: E731
    g, o = RoverBatch(mk()), oracle_lib.Oracle(mk(), "float")
    k, chunk, seen = 0, 1, 0
    while k < 6000:
        g.step(chunk); o.step(chunk); k += chunk
        sg = g.get_state(0)
        assert_states_equal(sg, o.get_state(0), 10, True, "%s step %d" % (kind, k))
        seen = max(seen, sg.n_contacts)
        chunk = 1 if k < 100 else 53
    assert seen >= 4
    g.close(); o.close()


def test_reference_terrain_batch_contact_counts_bit_exact():
    n = 128
    hm = reference_heightmap("rock_groove_large")
    mk = lambda: make_config(abi.PRESET_MT_RANDOM, n, heightfield=(hm, 6.0, 6.0, (2.0, 0.0, -0.3)), speedf=0.5, x_goal=4.0)  # noqa: E731
    g, o = RoverBatch(mk()), oracle_lib.Oracle(mk(), "float", threads=8)
    nc = torch.zeros(n, dtype=torch.int32, device="cuda")
    for k in range(40):
        g.step(50); o.step(50)
        g.observe(ncontacts=nc)
        assert np.array_equal(nc.cpu().numpy(), o.observe()[3]), "step %d" % ((k + 1) * 50)
    for e in range(0, n, 9):
        assert_states_equal(g.get_state(e), o.get_state(e), 10, True, "env %d" % e)
    g.close(); o.close()


@pytest.mark.parametrize("iters", [5, 8, 12, 16, 17, 30])
def test_other_sweep_counts_are_bit_exact(iters):
    """pgs_iters decides how many reorders a step has (one per 8 sweeps): none beyond the first (<= 8, no helper warp), one
    (9..16), two (17..24), and in-place shuffles past the third (> 24)."""
    hm = reference_heightmap("rock_groove")
    for kw in (dict(), dict(heightfield=(hm, 3.0, 3.0, (1.0, 0.0, -0.3)))):
        mk = lambda: make_config(abi.PRESET_MT_RANDOM, 4, pgs_iters=iters, speedf=0.4, **kw)  # noqa: E731
        g, o = RoverBatch(mk()), oracle_lib.Oracle(mk(), "float")
        for k in range(30):
            g.step(50); o.step(50)
            for e in range(4):
                assert_states_equal(g.get_state(e), o.get_state(e), nbodies(g.cfg), True, "iters %d step %d env %d" % (iters, (k + 1) * 50, e))
        g.close(); o.close()
