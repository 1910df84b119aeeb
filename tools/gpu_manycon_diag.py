"""Diagnostic: fp64 CUDA path vs oracle from identical many-contact states (curled, self-colliding fingers)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from mj_envs_b200.vec_env import AdroitVecEnv
from oracle.oracle import OracleSim
task = sys.argv[1] if len(sys.argv) > 1 else "relocate"
n = 128
env = AdroitVecEnv(f"{task}-v0", n, seed=3, fp64=True)
ora = OracleSim(task, n, seed=3)
env.reset(); ora.reset()
q, v, p = ora.get_state()
rng = np.random.default_rng(1)
q = q.copy(); q[:, :ora.nv - 6] += rng.uniform(-0.3, 0.9, (n, ora.nv - 6)) * float(os.environ.get("SC", "0.3"))
ora.set_state(q, np.zeros_like(v), p)
L = env.debug_layout()
rel = lambda x, y: np.abs(x - y).max(axis=-1) / np.maximum(1.0, np.abs(y).max(axis=-1))
for t in range(3):
    q, v, p = ora.get_state()
    env._set_raw_state(q, v, p); env.set_warmstart(ora.get_warmstart())
    a = rng.uniform(-1, 1, (n, env.nu)).astype(np.float32)
    ora.step(a); env.step(torch.from_numpy(a))
    _, ncon_o, nefc_o, _ = ora.solver()
    gi = np.zeros((n, L["GI_TOTAL"]), dtype=np.int32)
    env.abi.lib.adroit_debug_iscratch(env._h, gi.ctypes.data, None)
    gq, gv, _ = env._raw_state(); nq, nv_, _ = ora.get_state()
    eq, ev = rel(gq, nq), rel(gv, nv_)
    bad = np.where(eq > 1e-9)[0]
    print("step", t, "ncon max", ncon_o.max(), "n>16", int((ncon_o > 16).sum()), "bad envs", [(int(e), int(ncon_o[e]), int(gi[e, 0]), int(gi[e, 1]), int(gi[e, 2]), int(gi[e, 3]), float(eq[e]), float(ev[e])) for e in bad], "overflow", env.overflow())
