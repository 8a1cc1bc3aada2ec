"""ctypes wrapper of the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY."""
import ctypes as C
import os
import subprocess

import numpy as np

from tf_ode_b200 import abi

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB = None


def build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(_ROOT, "oracle")])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_ROOT, "oracle", "liboracle.so")
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.POINTER(abi.BatchConfig), C.c_int, C.c_int]
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_load_policy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.orc_step.argtypes = [C.c_void_p, C.c_int]
        L.orc_load_model_tree.argtypes = [C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 5 + [C.c_int] * 3
        L.orc_model_tree_predict.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_get_state.argtypes = [C.c_void_p, C.c_int, C.POINTER(abi.EnvState)]
        L.orc_set_state.argtypes = [C.c_void_p, C.c_int, C.POINTER(abi.EnvState)]
        L.orc_set_actions.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_set_speed.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_observe.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_drain_transitions.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.orc_reset.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.orc_last_rows.argtypes = [C.c_void_p, C.c_int]
        L.orc_max_pair_contacts.argtypes = [C.c_void_p]
        L.orc_contact_hist.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_lcp_solve.argtypes = [C.c_int] + [C.c_void_p] * 6
        L.orc_lcp_failures.argtypes = [C.c_void_p]
        L.orc_rand_ints.argtypes = [C.c_uint32, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_plane_space.argtypes = [C.c_void_p] * 3
        L.orc_philox.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]
        L.orc_biquad_coeffs.argtypes = [C.c_void_p]
        L.orc_biquad_run.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.orc_mlp_infer.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orc_sphere_plane.argtypes = [C.c_void_p, C.c_double, C.c_void_p, C.c_double, C.c_void_p]
        L.orc_sphere_box.argtypes = [C.c_void_p, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_cylinder_plane.argtypes = [C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_void_p, C.c_double, C.c_void_p]
        L.orc_box_plane.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_void_p]
        L.orc_sphere_hf.argtypes = [C.c_void_p, C.c_double, C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_double, C.c_void_p, C.c_void_p]
        L.orc_R_from_euler.argtypes = [C.c_double] * 3 + [C.c_void_p]
        L.orc_mass_box.argtypes = [C.c_double] * 4 + [C.c_void_p]
        L.orc_mass_cyl_z.argtypes = [C.c_double] * 3 + [C.c_void_p]
        _LIB = L
    return _LIB


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Oracle:
    """N independent rover environments stepped on the CPU. precision: 'float' | 'double'."""

    def __init__(self, cfg, precision="double", threads=1):
        self.L = lib()
        self.cfg = cfg
        self.n = cfg.n_envs
        self.h = self.L.orc_create(C.byref(cfg), 0 if precision == "float" else 1, threads)
        if not self.h:
            raise RuntimeError("orc_create failed")

    def close(self):
        if self.h:
            self.L.orc_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def load_policy(self, w):
        w = np.ascontiguousarray(w, dtype=np.float32)
        assert self.L.orc_load_policy(self.h, _p(w), w.size) == 0

    def load_model_trees(self, yaml_1, yaml_2, oblique=False, degree=1, interaction_only=False):
        """YAML text -> PyYAML dict (the library that wrote the reference's tree files) -> per-id entries for the oracle."""
        import yaml
        for which, text in enumerate((yaml_1, yaml_2)):
            d = yaml.safe_load(text)
            ids, term, off, cnt, par = [], [], [], [], []
            for k, v in d.items():
                p = v["model params"] if v["terminal"] else v["split params"]
                ids.append(int(k)); term.append(int(bool(v["terminal"]))); off.append(len(par)); cnt.append(len(p)); par.extend(float(x) for x in p)
            a = [np.ascontiguousarray(x, dtype=np.int32) for x in (ids, term, off, cnt)]
            par = np.ascontiguousarray(par, dtype=np.float64)
            rc = self.L.orc_load_model_tree(self.h, which, len(ids), _p(a[0]), _p(a[1]), _p(a[2]), _p(a[3]), _p(par), int(oblique), int(degree), int(interaction_only))
            if rc != 0:
                raise RuntimeError("oracle: model tree %d rejected" % (which + 1))

    def model_tree_predict(self, which, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.zeros(x.shape[0], np.float64)
        node = np.zeros(x.shape[0], np.int32)
        if self.L.orc_model_tree_predict(self.h, which, _p(x), x.shape[1], x.shape[0], _p(y), _p(node)) != 0:
            raise RuntimeError("oracle: model tree predict failed (dimension mismatch)")
        return y, node

    def step(self, n=1):
        self.L.orc_step(self.h, n)

    def get_state(self, env=0):
        s = abi.EnvState()
        self.L.orc_get_state(self.h, env, C.byref(s))
        return s

    def set_state(self, env, s):
        self.L.orc_set_state(self.h, env, C.byref(s))

    def set_actions(self, a):
        a = np.ascontiguousarray(a, dtype=np.float32).reshape(self.n, 2)
        self.L.orc_set_actions(self.h, _p(a))

    def set_speed(self, v):
        v = np.ascontiguousarray(v, dtype=np.float32).reshape(self.n)
        self.L.orc_set_speed(self.h, _p(v))

    def observe(self):
        obs = np.zeros((self.n, 17), np.float32)
        rew = np.zeros(self.n, np.float32)
        done = np.zeros(self.n, np.uint8)
        nc = np.zeros(self.n, np.int32)
        self.L.orc_observe(self.h, _p(obs), _p(rew), _p(done), _p(nc))
        return obs, rew, done, nc

    def drain_transitions(self, max_rows=1 << 16):
        out = np.zeros((max_rows, abi.TRANSITION_WIDTH), np.float32)
        n = self.L.orc_drain_transitions(self.h, _p(out), max_rows)
        return out[: min(n, max_rows)]

    def reset(self, ids=None):
        if ids is None:
            self.L.orc_reset(self.h, None, 0)
        else:
            ids = np.ascontiguousarray(ids, dtype=np.int32)
            self.L.orc_reset(self.h, _p(ids), ids.size)

    def contact_hist(self):
        """steps by number of GENERATED contacts (before the capacity cut), summed over environments; bin 128 = 128 or more."""
        h = np.zeros(129, np.uint64)
        self.L.orc_contact_hist(self.h, _p(h))
        return h

    def last_rows(self, env=0):
        return self.L.orc_last_rows(self.h, env)

    def bodies(self, env=0):
        """[12,13] pos, quat, lvel, avel."""
        s = self.get_state(env)
        return state_bodies(s)


def state_bodies(s):
    out = np.zeros((abi.NBODY, 13))
    out[:, 0:3] = np.array(s.pos)
    out[:, 3:7] = np.array(s.quat)
    out[:, 7:10] = np.array(s.lvel)
    out[:, 10:13] = np.array(s.avel)
    return out
