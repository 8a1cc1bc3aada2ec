"""The C++ host facade + rover_batch_mt driver (tf_ode_b200/host) against the Python mirror on the same scenes.  `-m gpu`."""
import os
import re
import subprocess

import numpy as np
import pytest

from tf_ode_b200 import PRESET_MT_TREE, RoverBatch, abi, make_config, reference_heightmap
from test_formats import encode_png

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DRIVER = os.path.join(ROOT, "tf_ode_b200", "host", "rover_batch_mt")
GETPOS = np.array([-0.1095, -0.023, 0.0075])   # Rover_1::GetPosition's point of the main body (src/rover_1.cc:248-253)


def _position(s):
    w, x, y, z = s.quat[0]
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                  [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                  [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
    return R @ GETPOS + np.array(s.pos[0])


def _run_driver(args):
    if not os.path.exists(DRIVER):
        pytest.fail("tf_ode_b200/host/rover_batch_mt is not built (python -c 'import __graft_entry__ as g; g.build()')")
    out = subprocess.run([DRIVER] + args, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-500:]
    m = re.search(r"envs (\d+) \| gpus (\d+) \| steps (\d+) \| finished episodes (\d+) \| success (\d+) \| mean x ([-\d.]+) \| mean y ([-+\d.]+)", out.stdout)
    assert m, out.stdout
    return out.stdout, dict(envs=int(m[1]), episodes=int(m[4]), success=int(m[5]), mean_x=float(m[6]), mean_y=float(m[7]))


def _python_stats(cfg, steps, trees=None):
    g = RoverBatch(cfg)
    if trees:
        g.load_model_trees(*trees)
    g.step(steps)
    st = [g.get_state(e) for e in range(cfg.n_envs)]
    g.close()
    p = np.array([_position(s) for s in st])
    return dict(episodes=sum(s.episodes_done for s in st), success=int((np.abs(p[:, 0]) >= cfg.x_goal).sum()),
                mean_x=float(p[:, 0].mean()), mean_y=float(p[:, 1].mean()))


def test_driver_with_the_reference_tree_file_prefix_matches_the_python_mirror():
    """argv[3] = tree-file prefix, as in src/scene_1_mt.cc:25-30 (P1.yaml / P2.yaml)."""
    n, steps = 48, 6000
    prefix = os.path.join(ROOT, "tests", "golden", "model_tree_")
    text, d = _run_driver(["nodisplay", "3", prefix, "0.1", "--envs", str(n), "--steps", str(steps)])
    assert "[Success]" in text or "[Failure]" in text   # the result line of src/scene_1_mt.cc:191-194
    y1, y2 = open(prefix + "1.yaml").read(), open(prefix + "2.yaml").read()
    cfg = make_config(PRESET_MT_TREE, n, auto_reset=0, step_yaw_deg=3.0, ic_start=1.1)
    want = _python_stats(cfg, steps, (y1, y2))
    assert d["envs"] == n and d["episodes"] == want["episodes"] and d["success"] == want["success"]
    assert d["mean_x"] == pytest.approx(want["mean_x"], abs=1e-3) and d["mean_y"] == pytest.approx(want["mean_y"], abs=1e-3)


def test_driver_on_a_png_heightmap_scene_matches_the_python_mirror(tmp_path):
    """--scene file.png: the HeightField image constructor of the facade (ode/heightfield.hh:42-59) end to end."""
    lv = np.load(os.path.join(ROOT, "tf_ode_b200", "data", "heightmap_levels.npz"))["rock_groove"]
    rgba = np.repeat(lv[..., None], 4, axis=2).astype(np.uint8); rgba[..., 3] = 255
    png = tmp_path / "heightmap_rock_groove.png"
    png.write_bytes(encode_png([rgba[y].tobytes() for y in range(lv.shape[0])], lv.shape[1], 8, 6, [0, 1, 2, 3, 4], idat_split=2))
    n, steps = 64, 4000
    _, d = _run_driver(["nodisplay", "0", "random", "0", "--envs", str(n), "--steps", str(steps), "--scene", str(png)])
    cfg = make_config(abi.PRESET_MT_RANDOM, n, auto_reset=0, heightfield=(reference_heightmap("rock_groove"), 3.0, 3.0, (1.0, 0.0, -0.3)))
    want = _python_stats(cfg, steps)
    assert d["episodes"] == want["episodes"] and d["success"] == want["success"]
    assert d["mean_x"] == pytest.approx(want["mean_x"], abs=1e-3) and d["mean_y"] == pytest.approx(want["mean_y"], abs=1e-3)
