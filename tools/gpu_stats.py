"""Workload statistics of a random-action rollout on the GPU box: contacts, constraint rows, Newton iterations, noslip sweeps.
Usage: python tools/gpu_stats.py --task hammer [--n 1024] [--steps 40]"""
import argparse, os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch  # noqa: E402
from mj_envs_b200.vec_env import AdroitVecEnv  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--task", default="hammer")
ap.add_argument("--n", type=int, default=1024)
ap.add_argument("--steps", type=int, default=40)
ap.add_argument("--fp64", action="store_true")
args = ap.parse_args()
env = AdroitVecEnv(f"{args.task}-v0", args.n, seed=0, auto_reset=True, fp64=args.fp64)
L = env.debug_layout()
env.reset()
g = torch.Generator(device=env.device); g.manual_seed(0)
acc = []
for t in range(args.steps):
    env.step(torch.rand((args.n, env.nu), device=env.device, generator=g) * 2 - 1)
    gi = np.zeros((args.n, L["GI_TOTAL"]), dtype=np.int32)
    env.abi.lib.adroit_debug_iscratch(env._h, gi.ctypes.data, None)
    acc.append(gi[:, :4].copy())
a = np.concatenate(acc)
for i, name in enumerate(["ncon", "nsimple", "newton iterations", "noslip sweeps"]):
    v = a[:, i]
    print(f"{args.task} {name:18s} mean {v.mean():6.2f}  p50 {np.percentile(v,50):4.0f} p90 {np.percentile(v,90):4.0f} p99 {np.percentile(v,99):4.0f} max {v.max():4d}   hist {np.bincount(v, minlength=8)[:12]} ... last bins {np.bincount(v, minlength=22)[18:22]}")
