"""Bit-level fingerprint of the CUDA path: runs every task for a number of random-action steps (fp32 and fp64 builds) and
prints a SHA-1 of the final qpos/qvel plus the summed rewards.  Two library builds whose results are meant to be
identical (a pure re-scheduling of the same arithmetic) must print the same lines:
  ADROIT_B200_LIB=build_ab/base.so python tools/gpu_fingerprint.py ; python tools/gpu_fingerprint.py"""
import hashlib, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np, torch
from mj_envs_b200.vec_env import AdroitVecEnv

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 60
for f64 in (False, True):
    for env_id, n in (("hammer-v0", 512), ("door-v0", 512), ("pen-v0", 512), ("relocate-v0", 512)):
        env = AdroitVecEnv(env_id, n, device=0, seed=3, auto_reset=True, fp64=f64)
        env.reset()
        g = torch.Generator(device="cuda:0"); g.manual_seed(7)
        tot = 0.0
        for i in range(steps if not f64 else steps // 3):
            a = torch.rand((n, env.nu), device="cuda:0", generator=g) * 2 - 1
            _, r, _, _ = env.step(a)
            tot += float(r.double().sum())
        q, v, _ = env._raw_state()
        h = hashlib.sha1(np.ascontiguousarray(q).tobytes() + np.ascontiguousarray(v).tobytes()).hexdigest()[:16]
        print(f"{env_id:12s} f64={int(f64)} {h} reward_sum={tot!r}")
        env.close()
