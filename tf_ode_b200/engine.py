"""Python host mirror of the reference's stepping interface, over the C ABI (include/tfode_batch.h).

`RoverBatch` plays the role of `ode::Environment` + `robot::Rover_1*` for N environments at once:
`step(n)` is n x { env.next_step(dt); robot.next_step(dt) } of the scene lambdas
(src/rover_training_1.cc:127-146).  There is no CPU implementation behind this class: if
libtfode_b200.so is missing or no B200 is visible, construction raises.
"""
import ctypes as C
import os

import numpy as np

from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TFODE_LIB") or os.path.join(_HERE, "csrc", "libtfode_b200.so")   # TFODE_LIB: developer override (A/B variants built by tools/build_variant.sh)
_LIB = None

EXPORTS = [
    "tfode_batch_config_default", "tfode_batch_create", "tfode_batch_destroy", "tfode_last_error",
    "tfode_batch_set_stream", "tfode_batch_reset", "tfode_batch_set_actions", "tfode_batch_set_speed",
    "tfode_batch_load_policy", "tfode_batch_step", "tfode_batch_observe", "tfode_batch_pack_rollout", "tfode_batch_get_bodies",
    "tfode_batch_drain_transitions", "tfode_batch_contact_stats", "tfode_batch_lcp_failures", "tfode_batch_rollout_host", "tfode_batch_get_state",
    "tfode_batch_set_state", "tfode_batch_launch_count", "tfode_batch_n_envs", "tfode_batch_sync",
    "tfode_policy_infer",
    "tfode_model_tree_parse", "tfode_batch_load_model_trees", "tfode_model_tree_infer", "tfode_batch_get_tree_nodes",
    "tfode_heightmap_load_png", "tfode_heightmap_decode_png", "tfode_free",
    "tfode_savedmodel_read_actor", "tfode_batch_load_policy_saved_model", "tfode_format_transition",
    "tfode_batch_step_dantzig_static",
]


class TfodeError(RuntimeError):
    pass


def load_library():
    """dlopen the C-ABI library. Raises if it was not built (run tf_ode_b200/csrc/build.sh)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise TfodeError("%s not found: build it with tf_ode_b200/csrc/build.sh (no CPU fallback exists)" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, i32, i64 = C.c_void_p, C.c_int, C.c_int64
    L.tfode_batch_config_default.argtypes = [C.POINTER(abi.BatchConfig), i32]
    L.tfode_batch_create.argtypes = [C.POINTER(abi.BatchConfig), C.POINTER(vp)]
    L.tfode_batch_destroy.argtypes = [vp]
    L.tfode_batch_destroy.restype = None
    L.tfode_last_error.argtypes = [vp]
    L.tfode_last_error.restype = C.c_char_p
    L.tfode_batch_set_stream.argtypes = [vp, vp]
    L.tfode_batch_reset.argtypes = [vp, vp, i32]
    L.tfode_batch_set_actions.argtypes = [vp, vp]
    L.tfode_batch_set_speed.argtypes = [vp, vp]
    L.tfode_batch_load_policy.argtypes = [vp, vp, C.c_size_t]
    L.tfode_batch_step.argtypes = [vp, i32]
    L.tfode_batch_observe.argtypes = [vp, vp, vp, vp, vp]
    L.tfode_batch_get_bodies.argtypes = [vp, vp]
    L.tfode_batch_pack_rollout.argtypes = [vp, vp]
    L.tfode_batch_drain_transitions.argtypes = [vp, vp, i32, C.POINTER(i32), C.POINTER(i32)]
    L.tfode_batch_contact_stats.argtypes = [vp, C.POINTER(i64), C.POINTER(i64), i32]
    L.tfode_batch_lcp_failures.argtypes = [vp, C.POINTER(i64)]
    L.tfode_batch_step_dantzig_static.argtypes = [vp, vp, i32, vp]
    L.tfode_batch_rollout_host.argtypes = [vp, vp, i32, vp, vp, vp]
    L.tfode_batch_get_state.argtypes = [vp, i32, C.POINTER(abi.EnvState)]
    L.tfode_batch_set_state.argtypes = [vp, i32, C.POINTER(abi.EnvState)]
    L.tfode_batch_launch_count.argtypes = [vp]
    L.tfode_batch_launch_count.restype = i64
    L.tfode_batch_n_envs.argtypes = [vp]
    L.tfode_batch_sync.argtypes = [vp]
    L.tfode_policy_infer.argtypes = [vp, vp, i32, vp]
    L.tfode_model_tree_parse.argtypes = [C.c_char_p, C.c_size_t, i32, C.POINTER(C.POINTER(abi.TreeNode)), C.POINTER(C.c_int32),
                                         C.POINTER(C.POINTER(C.c_double)), C.POINTER(C.c_int32)]
    L.tfode_batch_load_model_trees.argtypes = [vp, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t, i32, i32, i32]
    L.tfode_model_tree_infer.argtypes = [vp, vp, i32, vp, vp]
    L.tfode_batch_get_tree_nodes.argtypes = [vp, vp]
    L.tfode_savedmodel_read_actor.argtypes = [C.c_char_p, C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.c_size_t)]
    L.tfode_batch_load_policy_saved_model.argtypes = [vp, C.c_char_p]
    L.tfode_heightmap_load_png.argtypes = [C.c_char_p, C.c_double, C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    L.tfode_heightmap_decode_png.argtypes = [C.c_char_p, C.c_size_t, C.c_double, C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    L.tfode_format_transition.argtypes = [vp, i32, C.c_char_p, C.c_size_t]
    L.tfode_free.argtypes = [vp]
    L.tfode_free.restype = None
    _LIB = L
    return L


def parse_model_tree(yaml_text, oblique=False):
    """Linear_model_tree's constructor (scripts/ModelTree/cpp/model_tree.cc:11-45) on YAML text: the flattened tree as
    (nodes, params): nodes = list of dicts in the loader's depth-first order, params = float64 pool.  Raises TfodeError with
    the reference's runtime_error message on malformed trees."""
    L = load_library()
    if isinstance(yaml_text, str):
        yaml_text = yaml_text.encode()
    nodes, params = C.POINTER(abi.TreeNode)(), C.POINTER(C.c_double)()
    nn, npar = C.c_int32(0), C.c_int32(0)
    rc = L.tfode_model_tree_parse(yaml_text, len(yaml_text), int(oblique), C.byref(nodes), C.byref(nn), C.byref(params), C.byref(npar))
    if rc != 0:
        raise TfodeError(L.tfode_last_error(None).decode())
    try:
        out = [dict(id=nodes[i].id, terminal=bool(nodes[i].terminal), child=(nodes[i].child[0], nodes[i].child[1]),
                    param_off=nodes[i].param_off, n_params=nodes[i].n_params) for i in range(nn.value)]
        par = np.array([params[i] for i in range(npar.value)], dtype=np.float64)
    finally:
        L.tfode_free(nodes)
        L.tfode_free(params)
    return out, par


def load_heightmap_png(path_or_bytes, z_scale=0.3):
    """ode::HeightField's image constructor (ode/heightfield.hh:42-59): PNG -> float32 [nrow, ncol] heights in metres,
    row 0 = top row of the picture.  The reference hard-codes 0.3/255 whatever z_scale it is given."""
    L = load_library()
    h, nr, nc = C.POINTER(C.c_float)(), C.c_int32(0), C.c_int32(0)
    if isinstance(path_or_bytes, (bytes, bytearray)):
        rc = L.tfode_heightmap_decode_png(bytes(path_or_bytes), len(path_or_bytes), z_scale, C.byref(h), C.byref(nr), C.byref(nc))
    else:
        rc = L.tfode_heightmap_load_png(os.fsencode(path_or_bytes), z_scale, C.byref(h), C.byref(nr), C.byref(nc))
    if rc != 0:
        raise TfodeError(L.tfode_last_error(None).decode())
    try:
        return np.ctypeslib.as_array(h, shape=(nr.value, nc.value)).copy()
    finally:
        L.tfode_free(h)


def read_saved_model_actor(saved_model_dir):
    """TF_model's constructor (tf_cpp_binding.cc:14-66) without TensorFlow: the actor saved by
    `td3.actor.save( session_dir + '/actor' )` (scripts/rover_training_1.py:90,121) -> float32 blob W1 b1 W2 b2 W3 b3."""
    L = load_library()
    w, n = C.POINTER(C.c_float)(), C.c_size_t(0)
    if L.tfode_savedmodel_read_actor(os.fsencode(saved_model_dir), C.byref(w), C.byref(n)) != 0:
        raise TfodeError(L.tfode_last_error(None).decode())
    try:
        return np.ctypeslib.as_array(w, shape=(n.value,)).copy()
    finally:
        L.tfode_free(w)


def format_transitions(rows, kind=0):
    """PRINT_TRANSITIONS (kind 0, src/rover_1_tf.cc:125-136) / PRINT_STATE_AND_ACTIONS (kind 1, :190-195) text of experience rows
    ([m, abi.TRANSITION_WIDTH] float32, as drained): one line per row, exactly what the reference's printf calls emit."""
    L = load_library()
    rows = np.ascontiguousarray(rows, dtype=np.float32).reshape(-1, abi.TRANSITION_WIDTH)
    buf = C.create_string_buffer(1024)
    out = []
    for r in rows:
        n = L.tfode_format_transition(r.ctypes.data_as(C.c_void_p), kind, buf, len(buf))
        if n < 0:
            raise TfodeError("tfode_format_transition failed (%d)" % n)
        out.append(buf.raw[:n].decode())
    return "".join(out)


def _hp(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _dp(t):
    """device pointer of a torch tensor (or None)."""
    if t is None:
        return None
    assert t.is_cuda and t.is_contiguous()
    return C.c_void_p(t.data_ptr())


def glorot_policy(seed=7):
    """Glorot-uniform 17-512-512-2 actor with zero biases (no trained model ships), Keras [in,out] layout."""
    rng = np.random.RandomState(seed)
    out = []
    for fi, fo in ((17, 512), (512, 512), (512, 2)):
        lim = np.sqrt(6.0 / (fi + fo))
        out.append(rng.uniform(-lim, lim, size=(fi, fo)).astype(np.float32).ravel())
        out.append(np.zeros(fo, np.float32))
    w = np.concatenate(out)
    assert w.size == abi.POLICY_NFLOATS
    return w


class RoverBatch:
    def __init__(self, cfg, stream=None):
        self.L = load_library()
        self.cfg = cfg
        self.n = cfg.n_envs
        h = C.c_void_p()
        rc = self.L.tfode_batch_create(C.byref(cfg), C.byref(h))
        if rc != 0:
            raise TfodeError("tfode_batch_create failed (%d): %s" % (rc, self.L.tfode_last_error(None).decode()))
        self.h = h
        if stream is not None:
            self._ck(self.L.tfode_batch_set_stream(self.h, C.c_void_p(stream)))

    def _ck(self, rc):
        if rc != 0:
            raise TfodeError("tfode call failed (%d): %s" % (rc, self.L.tfode_last_error(self.h).decode()))

    def close(self):
        if getattr(self, "h", None):
            self.L.tfode_batch_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # --- the reference-facing calls -------------------------------------------------------------
    def step(self, n=1):
        self._ck(self.L.tfode_batch_step(self.h, n))

    def reset(self, ids=None):
        if ids is None:
            self._ck(self.L.tfode_batch_reset(self.h, None, 0))
        else:
            ids = np.ascontiguousarray(ids, dtype=np.int32)
            self._ck(self.L.tfode_batch_reset(self.h, _hp(ids), ids.size))

    def load_policy(self, w):
        w = np.ascontiguousarray(w, dtype=np.float32)
        self._ck(self.L.tfode_batch_load_policy(self.h, _hp(w), w.size))

    def load_policy_saved_model(self, saved_model_dir):
        """Rover_1_tf's `actor_path` (src/rover_1_tf.cc:24): the SavedModel directory itself."""
        self._ck(self.L.tfode_batch_load_policy_saved_model(self.h, os.fsencode(saved_model_dir)))

    def load_model_trees(self, yaml_1, yaml_2, oblique=False, degree=1, interaction_only=False):
        """Rover_1_mt's constructor (src/rover_1_mt.cc:13-27): YAML text of tree 1 (steering rate) and tree 2 (boggie torque)."""
        y1 = yaml_1.encode() if isinstance(yaml_1, str) else yaml_1
        y2 = yaml_2.encode() if isinstance(yaml_2, str) else yaml_2
        self._ck(self.L.tfode_batch_load_model_trees(self.h, y1, len(y1), y2, len(y2), int(oblique), int(degree), int(interaction_only)))

    def model_tree_infer(self, state_dev, act_dev, node_dev=None):
        """Rover_1_mt::InferAction batched: float64 [m,14] -> float64 [m,2] (+ int32 [m,2] terminal node ids)."""
        self._ck(self.L.tfode_model_tree_infer(self.h, _dp(state_dev), state_dev.shape[0], _dp(act_dev), _dp(node_dev)))

    def get_tree_nodes(self, out_dev):
        self._ck(self.L.tfode_batch_get_tree_nodes(self.h, _dp(out_dev)))

    def set_actions(self, actions_dev):
        self._ck(self.L.tfode_batch_set_actions(self.h, _dp(actions_dev)))

    def set_speed(self, speed_dev):
        self._ck(self.L.tfode_batch_set_speed(self.h, _dp(speed_dev)))

    def observe(self, obs=None, reward=None, done=None, ncontacts=None):
        self._ck(self.L.tfode_batch_observe(self.h, _dp(obs), _dp(reward), _dp(done), _dp(ncontacts)))

    def pack_rollout(self, rows_dev):
        """[n, abi.ROLLOUT_WIDTH] float32 device tensor: obs17 | reward | done | last episode return, steps | episodes | global id | 0."""
        self._ck(self.L.tfode_batch_pack_rollout(self.h, _dp(rows_dev)))

    def get_bodies(self, out_dev):
        self._ck(self.L.tfode_batch_get_bodies(self.h, _dp(out_dev)))

    def drain_transitions(self, out_dev, max_rows):
        """Experience rows appended since the last drain -> out_dev[:n]; returns n (rows written).  `last_dropped` = rows lost to
        the ring's transition_capacity since the previous drain.  Raises (nothing consumed) when the pending rows exceed max_rows."""
        n, dropped = C.c_int(0), C.c_int(0)
        self._ck(self.L.tfode_batch_drain_transitions(self.h, _dp(out_dev), max_rows, C.byref(n), C.byref(dropped)))
        self.last_dropped = dropped.value
        return n.value

    def contact_stats(self, reset=False):
        """(environment sub-steps in which more contacts were generated than max_contacts keeps, largest generated count)."""
        over, mx = C.c_int64(0), C.c_int64(0)
        self._ck(self.L.tfode_batch_contact_stats(self.h, C.byref(over), C.byref(mx), int(reset)))
        return over.value, mx.value

    def step_dantzig_static(self, actions_dev, n_substeps, bodies_out_dev):
        """tfode_batch_step_dantzig_static: set actions (or None), n dense-solver sub-steps on the library's static scratch, bodies out
        (or None) -- the call that allocates nothing; n_envs <= abi.NOALLOC_MAX_ENVS."""
        a = _dp(actions_dev) if actions_dev is not None else None
        o = _dp(bodies_out_dev) if bodies_out_dev is not None else None
        self._ck(self.L.tfode_batch_step_dantzig_static(self.h, a, int(n_substeps), o))

    def lcp_failures(self):
        n = C.c_int64(0)
        self._ck(self.L.tfode_batch_lcp_failures(self.h, C.byref(n)))
        return n.value

    def rollout_host(self, actions, n_substeps, obs, reward, done):
        """Host-buffer round trip (numpy arrays): H2D actions -> n steps -> D2H observations."""
        self._ck(self.L.tfode_batch_rollout_host(self.h, _hp(actions), n_substeps, _hp(obs), _hp(reward), _hp(done)))

    def policy_infer(self, obs_scaled_dev, act_dev):
        self._ck(self.L.tfode_policy_infer(self.h, _dp(obs_scaled_dev), obs_scaled_dev.shape[0], _dp(act_dev)))

    def get_state(self, env=0):
        s = abi.EnvState()
        self._ck(self.L.tfode_batch_get_state(self.h, env, C.byref(s)))
        return s

    def set_state(self, env, s):
        self._ck(self.L.tfode_batch_set_state(self.h, env, C.byref(s)))

    def sync(self):
        self._ck(self.L.tfode_batch_sync(self.h))

    @property
    def launch_count(self):
        return self.L.tfode_batch_launch_count(self.h)
