"""Generates tests/golden/*.npy from the REFERENCE'S OWN code, run in this container.

  biquad_x.npy / biquad_y.npy : the reference's filters::LP_second_order_bilinear<double> (Filters/cpp/filters.cc,
      compiled in place by `make -C oracle ref` into oracle/_ref/libref_filters.so), constructed exactly as
      src/rover_1.cc:244 does (Te = 0.001f, w0 = 4 pi, Q = 0.5, in place on one double), driven by a seeded
      input made of a step, a chirp and noise.
  heightmap_levels.npz : the red channel (what ode/heightfield.hh:54 reads through osg::Image::data(c, r)) of the
      reference's env_data/heightmap_rock_{step,groove,groove_large}.png, decoded here with PIL, row 0 = top row of the
      picture (= row 0 of the array handed to dGeomHeightfieldDataBuildDouble after heightfield.hh's row flip of OSG's
      bottom-up image).  uint8 levels; height = level * 0.3 / 255.
  model_tree_{1,2}.yaml / model_tree_io.npz : two straight linear model trees fitted by the reference's OWN
      scripts/ModelTree/model_tree.py (imported from /root/reference; its `cma` import, used only by oblique-split
      search, is stubbed) on seeded synthetic 14-feature data (the Rover_1_mt::GetState layout, src/rover_1_mt.cc:30-52),
      saved with its save_tree_params() (PyYAML), plus its predict(X, return_node_id=True) on seeded inputs.
      model_tree_poly2.yaml: the same with model='polynomial', degree 2 (Polynomial_model_tree in the C++ loader).
Run from the repo root:  make -C oracle ref && python tools/make_golden.py
/root/reference is needed only here; the tests read the committed .npy files.
"""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    ref = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_filters.so"))
    ref.ref_biquad_run.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    rng = np.random.RandomState(20260411)
    n = 4000
    t = np.arange(n) * 0.001
    x = np.where(t > 0.2, 100.0, 0.0) + 30.0 * np.sin(2 * np.pi * (1 + 20 * t) * t) + 5.0 * rng.randn(n)
    x = np.ascontiguousarray(x, dtype=np.float64)
    y = np.zeros_like(x)
    ref.ref_biquad_run(x.ctypes.data_as(C.c_void_p), n, y.ctypes.data_as(C.c_void_p))
    os.makedirs(os.path.join(ROOT, "tests", "golden"), exist_ok=True)
    np.save(os.path.join(ROOT, "tests", "golden", "biquad_x.npy"), x)
    np.save(os.path.join(ROOT, "tests", "golden", "biquad_y.npy"), y)
    print("biquad golden: n=%d, y[-1]=%.17g" % (n, y[-1]))


def heightmaps():
    from PIL import Image
    out = {}
    for k in ("rock_step", "rock_groove", "rock_groove_large"):
        a = np.array(Image.open("/root/reference/env_data/heightmap_%s.png" % k))
        assert a.ndim == 3 and a.dtype == np.uint8
        out[k] = np.ascontiguousarray(a[..., 0])
        print("heightmap %s: %s levels %d..%d" % (k, out[k].shape, out[k].min(), out[k].max()))
    np.savez_compressed(os.path.join(ROOT, "tf_ode_b200", "data", "heightmap_levels.npz"), **out)   # package data: the terrains of BASELINE configs 2 and 5


def model_trees():
    import sys
    import types
    sys.modules.setdefault("cma", types.ModuleType("cma"))  # only CMA_search (oblique splits) uses it
    sys.path.insert(0, "/root/reference/scripts/ModelTree")
    import model_tree as ref_mt

    rng = np.random.RandomState(20260411)
    scale = np.array([90, 45, 25, 25, 45, 100, 100, 100, 30, 30, 30, 30, 30, 30], dtype=np.float64)
    X = rng.uniform(-1, 1, size=(3000, 14)) * scale
    # piecewise-linear targets in the range of the two commands (steering rate deg/s, boggie torque N.m)
    y1 = np.where(X[:, 1] >= 5.0, -8.0 + 0.1 * X[:, 0], np.where(X[:, 3] >= -3.0, 6.0 - 0.2 * X[:, 2], 0.05 * X[:, 7])) + 0.01 * rng.randn(3000)
    y2 = np.where(X[:, 4] >= 0.0, 12.0 + 0.1 * X[:, 5] - 0.3 * X[:, 3], -10.0 + 0.2 * X[:, 12]) + 0.01 * rng.randn(3000)
    Xt = rng.uniform(-1, 1, size=(512, 14)) * scale
    gd = os.path.join(ROOT, "tests", "golden")
    io = {"X": Xt}
    for name, y, kw in (("1", y1, dict(model="linear")), ("2", y2, dict(model="linear")),
                        ("poly2", y2 + 0.01 * X[:, 0] * X[:, 1], dict(model="polynomial", degree=2))):
        t = ref_mt.Model_tree(oblique=False, max_depth=3, node_min_samples=50, search_grid=10, **kw)
        t.fit(X, y, verbose=0)
        t.save_tree_params(os.path.join(gd, "model_tree_" + name))
        pred, nodes = t.predict(Xt, return_node_id=True)
        io["y_" + name] = np.asarray(pred, dtype=np.float64)
        io["node_" + name] = np.asarray(nodes, dtype=np.int32)
        print("model tree %s: %d nodes reached, y[:3] = %s" % (name, len(set(nodes)), io["y_" + name][:3]))
    np.savez_compressed(os.path.join(gd, "model_tree_io.npz"), **io)


if __name__ == "__main__":
    main()
    heightmaps()
    model_trees()
