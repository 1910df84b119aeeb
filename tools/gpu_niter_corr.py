"""How well does the Newton iteration count of substep t predict substep t+1 of the same env?  (hammer at frame_skip 1, so that every
env.step() is one substep and GI_NITER can be read after each.)  Prints the lag-1 transition table and E[max of 5] for random blocks
against blocks sorted by the previous count."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch  # noqa: E402
from mj_envs_b200.vec_env import AdroitVecEnv  # noqa: E402
from mj_envs_b200.mjcf import load_compiled, Model  # noqa: E402
task = sys.argv[1] if len(sys.argv) > 1 else "hammer"
m = load_compiled(task)
m1 = Model(task=m.task, names=m.names, A=dict(m.A)); m1.A["sizes"] = m1.A["sizes"].copy(); m1.A["sizes"][9] = 1
n = 2048
env = AdroitVecEnv(f"{task}-v0", n, seed=0, model=m1, auto_reset=True)
L = env.debug_layout()
env.reset()
g = torch.Generator(device=env.device); g.manual_seed(0)
its, ncons, nsimple = [], [], []
a = torch.rand((n, env.nu), device=env.device, generator=g) * 2 - 1
for t in range(300):
    if t % 5 == 0: a = torch.rand((n, env.nu), device=env.device, generator=g) * 2 - 1   # the action holds for 5 substeps, as in a step
    env.step(a)
    gi = np.zeros((n, L["GI_TOTAL"]), dtype=np.int32)
    env.abi.lib.adroit_debug_iscratch(env._h, gi.ctypes.data, None)
    its.append(gi[:, 2].copy()); ncons.append(gi[:, 0].copy()); nsimple.append(gi[:, 1].copy())
its = np.array(its[50:]); ncons = np.array(ncons[50:]); nsimple = np.array(nsimple[50:])
x, y = its[:-1].ravel(), its[1:].ravel()
print("mean iterations", its.mean(), "corr lag1", np.corrcoef(x, y)[0, 1])
T = np.zeros((9, 9), int)
for i, j in zip(np.minimum(x, 8), np.minimum(y, 8)): T[i, j] += 1
print("transition counts (rows: previous, cols: next)\n", T)
rng = np.random.default_rng(0)
def emax(order_key, lag=1):
    tot = 0; cnt = 0
    for t in range(1, its.shape[0]):
        idx = np.argsort(order_key[t - lag], kind="stable") if order_key is not None else rng.permutation(n)
        v = its[t][idx][: n // 5 * 5].reshape(-1, 5)
        tot += v.max(axis=1).sum(); cnt += v.shape[0]
    return tot / cnt
print("E[max of 5] random blocks", emax(None), " sorted by previous count", emax(its), " mean", its.mean())
# predictors known BEFORE the solve kernel of substep t runs: the contact count the collide kernel of the same substep left (lag 0)
print("corr(iterations, ncon same substep)", np.corrcoef(its.ravel(), ncons.ravel())[0, 1],
      " corr(iterations, simple rows)", np.corrcoef(its.ravel(), nsimple.ravel())[0, 1])
print("E[max of 5] sorted by ncon of the same substep", emax(ncons, 0), " by ncon then previous count", 
      emax(np.concatenate([ncons[:1] * 100, ncons[1:] * 100 + its[:-1]]), 0), " oracle (sorted by own count)", emax(its, 0))
for c in range(0, 12):
    sel = ncons == c
    if sel.sum(): print(f"ncon {c:2d}: share {sel.mean():.3f} mean iterations {its[sel].mean():.2f}")
sel = ncons >= 12
if sel.sum(): print(f"ncon>=12: share {sel.mean():.3f} mean iterations {its[sel].mean():.2f}")
