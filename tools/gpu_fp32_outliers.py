"""fp32 CUDA path vs fp64 oracle, single steps from identical states at frame_skip 1 (the loop of
tests/test_gpu_round2.py::test_fp32_tolerance_holds_wherever_the_constraint_set_agrees), printing the numbers the test only bounds:
env-steps with agreeing / differing constraint sets, the share of agreeing ones inside 1e-4, the worst agreeing error, the outliers."""
import os, sys
import numpy as np
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import torch  # noqa: E402
from mj_envs_b200.vec_env import AdroitVecEnv  # noqa: E402
from mj_envs_b200.mjcf import load_compiled, Model  # noqa: E402
from oracle.oracle import OracleSim  # noqa: E402
from test_gpu_round2 import _constraint_sets, _rel  # noqa: E402
for task in (sys.argv[1:] or ["hammer", "door", "pen", "relocate"]):
    n, steps = 96, 60
    m = load_compiled(task)
    m1 = Model(task=m.task, names=m.names, A=dict(m.A)); m1.A["sizes"] = m1.A["sizes"].copy(); m1.A["sizes"][9] = 1
    env = AdroitVecEnv(f"{task}-v0", n, seed=3, model=m1); ora = OracleSim(m1, n, seed=3)
    env.reset(); ora.reset()
    rng = np.random.default_rng(3)
    agree = differ = within = 0; worst = 0.0; out = []; worst_differ = 0.0
    for t in range(steps):
        q, v, p = ora.get_state()
        env._set_raw_state(q, v, p); env.set_warmstart(ora.get_warmstart())
        a = rng.uniform(-1, 1, (n, env.nu)).astype(np.float32)
        if t % 8 < 4: a[:] = rng.uniform(-1, 1, (1, env.nu)).astype(np.float32)
        ora.step(a); env.step(torch.from_numpy(a))
        same = _constraint_sets(env, ora)
        gq, gv, _ = env._raw_state(); nq, nv_, _ = ora.get_state()
        e = np.maximum(_rel(gq, nq), _rel(gv, nv_))
        agree += int(same.sum()); differ += int((~same).sum()); within += int((same & (e < 1e-4)).sum())
        if same.any(): worst = max(worst, float(e[same].max()))
        if (~same).any(): worst_differ = max(worst_differ, float(e[~same].max()))
        out += [float(e[i]) for i in np.nonzero(same & (e >= 1e-4))[0]]
    print(f"{task:9s} agreeing {agree} differing {differ} ({100 * differ / (agree + differ):.2f} %)  agreeing within 1e-4: {within} ({100 * within / max(agree, 1):.3f} %)  "
          f"worst agreeing {worst:.2e}  outliers {sorted(out)[-6:]}  worst differing {worst_differ:.2e}")
    env.close()
