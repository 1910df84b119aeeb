"""numpy front-end of the fp64 CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(_HERE, ".."))
from mj_envs_b200._cabi import CABI  # noqa: E402
from mj_envs_b200.mjcf import Model, load_compiled  # noqa: E402

_LIB = os.path.join(_HERE, "liboracle.so")


def build(force=False):
    if force or not os.path.exists(_LIB):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _LIB


_abi = None


def abi():
    global _abi
    if _abi is None:
        _abi = CABI(build(), "oracle")
        L = _abi.lib
        L.oracle_debug_dynamics.restype = C.c_int
        L.oracle_debug_dynamics.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 5
        L.oracle_debug_contacts.restype = C.c_int
        L.oracle_debug_contacts.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.oracle_debug_collide.restype = C.c_int
        L.oracle_debug_collide.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_double, C.c_void_p, C.c_int]
    return _abi


def _p(a):
    return None if a is None else a.ctypes.data


class OracleSim:
    def __init__(self, task_or_model, n_envs=1, seed=0, env_id_offset=0, flags=0):
        self.model = task_or_model if isinstance(task_or_model, Model) else load_compiled(task_or_model)
        self.abi = abi()
        blob = self.model.to_blob()
        h = C.c_void_p()
        self.abi.check(self.abi.create(blob, len(blob), n_envs, 0, seed, env_id_offset, flags, C.byref(h)))
        self.h = h
        self.n = n_envs
        sz = np.zeros(8, dtype=np.int32)
        self.abi.sizes(h, _p(sz))
        self.nv, self.nu, self.obs_dim, self.npar, self.horizon, self.frame_skip, self.nbody, self.nsite = map(int, sz)

    def __del__(self):
        if getattr(self, "h", None):
            self.abi.destroy(self.h)
            self.h = None

    def reset(self, mask=None):
        obs = np.zeros((self.n, self.obs_dim), dtype=np.float32)
        m = None if mask is None else np.ascontiguousarray(mask, dtype=np.uint8)
        self.abi.check(self.abi.reset(self.h, _p(m), _p(obs), None))
        return obs

    def step(self, action):
        a = np.ascontiguousarray(action, dtype=np.float32).reshape(self.n, self.nu)
        obs = np.zeros((self.n, self.obs_dim), dtype=np.float32)
        rew = np.zeros(self.n, dtype=np.float32)
        done = np.zeros(self.n, dtype=np.uint8)
        goal = np.zeros(self.n, dtype=np.uint8)
        self.abi.check(self.abi.step(self.h, _p(a), _p(obs), _p(rew), _p(done), _p(goal), None))
        return obs, rew, done, goal

    def get_obs(self):
        obs = np.zeros((self.n, self.obs_dim), dtype=np.float32)
        self.abi.check(self.abi.get_obs(self.h, _p(obs), None))
        return obs

    def get_state(self):
        q = np.zeros((self.n, self.nv)); v = np.zeros((self.n, self.nv)); p = np.zeros((self.n, self.npar))
        self.abi.check(self.abi.get_state(self.h, _p(q), _p(v), _p(p), None))
        return q, v, p

    def set_state(self, qpos=None, qvel=None, par=None):
        c = lambda a, w: None if a is None else np.ascontiguousarray(a, dtype=np.float64).reshape(self.n, w)
        q, v, p = c(qpos, self.nv), c(qvel, self.nv), c(par, self.npar)
        self.abi.check(self.abi.set_state(self.h, _p(q), _p(v), _p(p), None))

    def set_elapsed_steps(self, t):
        t = np.ascontiguousarray(t, dtype=np.int32).reshape(self.n)
        self.abi.check(self.abi.set_elapsed_steps(self.h, _p(t), None))

    def get_warmstart(self):
        w = np.zeros((self.n, self.nv))
        self.abi.check(self.abi.get_warmstart(self.h, _p(w), None))
        return w

    def set_warmstart(self, w):
        w = np.ascontiguousarray(w, dtype=np.float64).reshape(self.n, self.nv)
        self.abi.check(self.abi.set_warmstart(self.h, _p(w), None))

    def frames(self):
        bx = np.zeros((self.n, self.nbody, 3)); bq = np.zeros((self.n, self.nbody, 4)); sx = np.zeros((self.n, self.nsite, 3))
        self.abi.check(self.abi.get_frames(self.h, _p(bx), _p(bq), _p(sx), None))
        return bx, bq, sx

    def solver(self):
        qacc = np.zeros((self.n, self.nv)); ncon = np.zeros(self.n, dtype=np.int32); nefc = np.zeros(self.n, dtype=np.int32)
        touch = np.zeros(self.n)
        self.abi.check(self.abi.get_solver(self.h, _p(qacc), _p(ncon), _p(nefc), _p(touch), None))
        return qacc, ncon, nefc, touch

    def reduce_episode(self, flush=False):
        out = np.zeros(4)
        self.abi.check(self.abi.reduce_episode(self.h, int(flush), _p(out), None))
        return out

    # ---- oracle-only introspection
    def dynamics(self, env=0):
        nv = self.nv
        M = np.zeros((nv, nv)); bias = np.zeros(nv); smooth = np.zeros(nv); qas = np.zeros(nv); qfc = np.zeros(nv)
        self.abi.lib.oracle_debug_dynamics(self.h, env, _p(M), _p(bias), _p(smooth), _p(qas), _p(qfc))
        return dict(M=M, qfrc_bias=bias, qfrc_smooth=smooth, qacc_smooth=qas, qfrc_constraint=qfc)

    def contacts(self, env=0):
        out = np.zeros((100, 16))
        n = self.abi.lib.oracle_debug_contacts(self.h, env, _p(out), 100)
        return out[:n]


def collide(type1, pos1, mat1, size1, type2, pos2, mat2, size2, margin=0.0):
    a = abi()
    f = lambda x, n: np.ascontiguousarray(np.asarray(x, dtype=np.float64).reshape(-1)[:n])
    out = np.zeros((16, 16))
    s1 = np.zeros(3); s1[:len(size1)] = size1
    s2 = np.zeros(3); s2[:len(size2)] = size2
    p1, m1, p2, m2 = f(pos1, 3), f(mat1, 9), f(pos2, 3), f(mat2, 9)
    n = a.lib.oracle_debug_collide(type1, _p(p1), _p(m1), _p(s1), type2, _p(p2), _p(m2), _p(s2), margin, _p(out), 16)
    return out[:n]


def efc(sim: OracleSim, env=0):
    """Constraint rows of one env (re-runs forward): returns J [nefc,nv] and a dict of per-row columns."""
    L = abi().lib
    L.oracle_debug_efc.restype = C.c_int
    L.oracle_debug_efc.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
    J = np.zeros((500, sim.nv)); cols = np.zeros((500, 8))
    n = L.oracle_debug_efc(sim.h, env, _p(J), _p(cols), 500)
    keys = ["type", "aref", "R", "floss", "force", "pos", "margin", "diagApprox"]
    return J[:n], {k: cols[:n, i] for i, k in enumerate(keys)}
