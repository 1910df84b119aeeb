"""Soak run: long random-action rollouts with auto-reset on the product path; checks finiteness, capacity overflow and that the
episode bookkeeping adds up.  Usage: python tools/gpu_soak.py [steps] [envs]"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch  # noqa: E402
from mj_envs_b200.vec_env import AdroitVecEnv  # noqa: E402
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 600
n = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
for task in ("hammer", "door", "pen", "relocate"):
    env = AdroitVecEnv(f"{task}-v0", n, seed=7, auto_reset=True, reset_on_done=(task == "pen"))
    obs, _ = env.reset()
    g = torch.Generator(device=env.device); g.manual_seed(1)
    bad = 0; dones = 0; rmin, rmax = 1e30, -1e30
    t0 = time.time()
    for t in range(steps):
        scale = 1.0 if (t // 50) % 2 == 0 else 3.0          # every other block of 50 steps: actions far outside [-1, 1] (clipped by the action map)
        obs, r, d, info = env.step((torch.rand((n, env.nu), device=env.device, generator=g) * 2 - 1) * scale)
        bad += int((~torch.isfinite(obs)).any(dim=1).sum()) + int((~torch.isfinite(r)).sum())
        dones += int(d.sum()); rmin = min(rmin, float(r.min())); rmax = max(rmax, float(r.max()))
    torch.cuda.synchronize()
    stats = env.reduce_episode(flush_running=False, all_reduce=False)
    ov = env.overflow()
    print(f"{task:9s} {steps} steps x {n} envs in {time.time() - t0:.1f} s: non-finite rows {bad}, done flags {dones}, finished episodes {int(stats[3])}, "
          f"reward range [{rmin:.2f}, {rmax:.2f}], overflow {ov.tolist()}")
    assert bad == 0 and ov.sum() == 0
    env.close()
print("soak ok")
