"""Stage-by-stage GPU-vs-oracle diagnostic (run on the GPU box): prints the max abs error of every pipeline stage.
Usage: python tools/gpu_diag.py [--fp32] [--tasks hammer,door] [--n 32] [--steps 20]"""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch  # noqa: E402
from mj_envs_b200.vec_env import AdroitVecEnv  # noqa: E402
from oracle.oracle import OracleSim  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--fp32", action="store_true")
ap.add_argument("--tasks", default="hammer,door,pen,relocate")
ap.add_argument("--n", type=int, default=32)
ap.add_argument("--steps", type=int, default=30)
args = ap.parse_args()


def tri_to_full(p, nv):
    M = np.zeros((nv, nv))
    k = 0
    for i in range(nv):
        for j in range(i + 1):
            M[i, j] = M[j, i] = p[k]
            k += 1
    return M


def err(a, b):
    return float(np.abs(np.asarray(a, dtype=np.float64) - np.asarray(b, dtype=np.float64)).max())


for task in args.tasks.split(","):
    n = args.n
    env = AdroitVecEnv(f"{task}-v0", n, seed=7, fp64=not args.fp32)
    ora = OracleSim(task, n, seed=7)
    L = env.debug_layout()
    o_obs = ora.reset()
    g_obs, _ = env.reset()
    torch.cuda.synchronize()
    print(f"== {task} ({'fp32' if args.fp32 else 'fp64'}) nv={env.nv}")
    print("reset: obs err", err(g_obs.cpu().numpy(), o_obs), " par err", err(env._raw_state()[2], ora.get_state()[2]))
    rng = np.random.default_rng(0)
    m = ora.model
    lo, hi = m.jnt_range[:, 0], m.jnt_range[:, 1]
    q = np.where(m.jnt_limited > 0, rng.uniform(lo, hi, (n, m.nv)), rng.uniform(-0.2, 0.2, (n, m.nv)))
    v = rng.normal(0, 0.5, (n, m.nv))
    par = ora.get_state()[2]
    ora.set_state(q, v, par); ora.set_warmstart(np.zeros((n, m.nv)))
    env._set_raw_state(q, v, par); env.set_warmstart(np.zeros((n, m.nv)))
    ora.set_state(q, v, par); env._set_raw_state(q, v, par)
    bx, bq, sx = env.frames()
    obx, obq, osx = ora.frames()
    print("random state: xpos err", err(bx, obx), " site err", err(sx, osx),
          " xquat err", float(np.minimum(np.abs(bq - obq).max(axis=2), np.abs(bq + obq).max(axis=2)).max()))
    nv = m.nv
    Mg = env.debug_scratch(L["G_DBG_M"], nv * (nv + 1) // 2)
    bias = env.debug_scratch(L["G_DBG_BIAS"], nv)
    qfs = env.debug_scratch(L["G_DBG_QFS"], nv)
    qas = env.debug_scratch(L["G_DBG_QAS"], nv)
    gq, gncon, gnefc, gtouch = env.solver()
    oq, oncon, onefc, otouch = ora.solver()
    eM = ebias = eqfs = eqas = 0
    for e in range(n):
        d = ora.dynamics(e)
        eM = max(eM, err(tri_to_full(Mg[e], nv), d["M"]))
        ebias = max(ebias, err(bias[e], d["qfrc_bias"]))
        eqfs = max(eqfs, err(qfs[e], d["qfrc_smooth"]))
        eqas = max(eqas, err(qas[e], d["qacc_smooth"]) / max(1.0, np.abs(d["qacc_smooth"]).max()))
    print(f"  M err {eM:.3e}  bias err {ebias:.3e}  qfrc_smooth err {eqfs:.3e}  qacc_smooth rel err {eqas:.3e}")
    print("  ncon gpu", gncon[:12], "oracle", oncon[:12], " mismatches", int((gncon != oncon).sum()))
    print("  nefc mismatches", int((gnefc != onefc).sum()), " qacc rel err per env (max)",
          float((np.abs(gq - oq).max(axis=1) / np.maximum(1, np.abs(oq).max(axis=1))).max()))
    # contacts of env with most contacts
    e = int(np.argmax(oncon))
    oc = ora.contacts(e)
    gc = env.debug_scratch(L["G_CON"], 16 * L["MAXCON"])[e].reshape(-1, 16)[: gncon[e]]
    if len(oc) and len(gc) == len(oc):
        print("  contacts env", e, "dist err", err(gc[:, 0], oc[:, 0]), "pos err", err(gc[:, 1:4], oc[:, 1:4]), "normal err", err(gc[:, 4:7], oc[:, 4:7]))
    # single-step parity from identical states along an oracle trajectory
    ora.reset(); env.reset()
    worst = dict(qpos=0.0, qvel=0.0, rew=0.0, obs=0.0)
    bad_ncon = 0
    t0 = time.time()
    for t in range(args.steps):
        oq_, ov_, op_ = ora.get_state()
        ow_ = ora.get_warmstart()
        env._set_raw_state(oq_, ov_, op_); env.set_warmstart(ow_)
        a = rng.uniform(-1, 1, (n, env.nu)).astype(np.float32)
        oobs, orew, odone, ogoal = ora.step(a)
        gobs, grew, gdone, ginfo = env.step(torch.from_numpy(a))
        gq_, gv_, _ = env._raw_state()
        nq_, nv_, _ = ora.get_state()
        sq = np.maximum(1.0, np.abs(nq_).max(axis=1, keepdims=True)); sv = np.maximum(1.0, np.abs(nv_).max(axis=1, keepdims=True))
        worst["qpos"] = max(worst["qpos"], float((np.abs(gq_ - nq_) / sq).max()))
        worst["qvel"] = max(worst["qvel"], float((np.abs(gv_ - nv_) / sv).max()))
        worst["rew"] = max(worst["rew"], float((np.abs(grew.cpu().numpy() - orew) / np.maximum(1, np.abs(orew))).max()))
        worst["obs"] = max(worst["obs"], err(gobs.cpu().numpy(), oobs))
    print(f"  single-step parity over {args.steps} steps x {n} envs (rel. to max(1,|x|_inf)):", {k: f"{v_:.2e}" for k, v_ in worst.items()},
          " overflow", env.overflow(), f" [{time.time()-t0:.1f}s]")
    # free-running comparison (no state sync)
    ora.reset(); env.reset()
    for t in range(args.steps):
        a = rng.uniform(-1, 1, (n, env.nu)).astype(np.float32)
        oobs, orew, _, _ = ora.step(a)
        gobs, grew, _, _ = env.step(torch.from_numpy(a))
        if t in (0, 4, 9, 19, args.steps - 1):
            print(f"    free-run t={t+1}: obs err {err(gobs.cpu().numpy(), oobs):.2e} reward err {err(grew.cpu().numpy(), orew):.2e}")
    env.close()
