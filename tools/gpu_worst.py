"""Locates the worst single-step fp32-vs-oracle discrepancy and prints what differs at that state (GPU box only).
Usage: python tools/gpu_worst.py --task pen [--n 64] [--steps 60] [--fp64]"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch  # noqa: E402
from mj_envs_b200.vec_env import AdroitVecEnv  # noqa: E402
from oracle.oracle import OracleSim  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--task", default="pen")
ap.add_argument("--n", type=int, default=64)
ap.add_argument("--steps", type=int, default=60)
ap.add_argument("--fp64", action="store_true")
ap.add_argument("--frame-skip", type=int, default=1, help="substeps per compared step (1 isolates a single mj_step)")
args = ap.parse_args()

n = args.n
env = AdroitVecEnv(f"{args.task}-v0", n, seed=5, fp64=args.fp64, frame_skip=args.frame_skip)
model = env.model
ora = OracleSim(model, n, seed=5)
L = env.debug_layout()
ora.reset(); env.reset()
rng = np.random.default_rng(1)
records = []
import collections
cls = collections.Counter(); nbig = 0; ntot = 0
worst = (0.0, None)
for t in range(args.steps):
    q, v, p = ora.get_state()
    w = ora.get_warmstart()
    env._set_raw_state(q, v, p); env.set_warmstart(w)
    ora.set_state(q, v, p)                       # re-run the oracle's forward at this state (ctrl persists)
    oqa, oncon, onefc, _ = ora.solver()
    gqa, gncon, gnefc, _ = env.solver()          # forward at the same state and ctrl
    eacc = np.abs(gqa - oqa).max(axis=1) / np.maximum(1.0, np.abs(oqa).max(axis=1))
    gcon = env.debug_scratch(L["G_CON"], 16 * L["MAXCON"]).reshape(n, -1, 16)
    gi_ = np.zeros((n, L["GI_TOTAL"]), dtype=np.int32)
    env.abi.lib.adroit_debug_iscratch(env._h, gi_.ctypes.data, None)
    for e in np.nonzero(eacc > 1e-3)[0]:
        if gncon[e] != oncon[e]:
            cls["ncon differs"] += 1
            continue
        oc = ora.contacts(e)
        dn = np.abs(gcon[e, :gncon[e], 4:7] - oc[:, 4:7]).max(axis=1) if len(oc) else np.zeros(0)
        dd = np.abs(gcon[e, :gncon[e], 0] - oc[:, 0]) if len(oc) else np.zeros(0)
        if len(dn) and dn.max() > 1e-3:
            k = int(np.argmax(dn))
            g = gi_[e, L["GI_CON"] + 4 * k: L["GI_CON"] + 4 * k + 2]
            cls["normal differs, types %s" % (tuple(int(model.A["geom_type"][x]) for x in g),)] += 1
        elif len(dd) and dd.max() > 1e-5:
            cls["dist differs"] += 1
        else:
            cls["solver (same contacts) ncon=%d iters=%d" % (gncon[e], gi_[e, 2])] += 1
    nbig += int((eacc > 1e-3).sum()); ntot += n
    a = rng.uniform(-1, 1, (n, env.nu)).astype(np.float32)
    if t % 10 < 5:
        a[:] = a[:, :1] * 0 + rng.uniform(-1, 1, (1, env.nu)).astype(np.float32)   # correlated grasping motions
    ora.step(a)
    env.step(torch.from_numpy(a))
    gq, gv, _ = env._raw_state()
    nq, nv_, _ = ora.get_state()
    ev = np.abs(gv - nv_).max(axis=1) / np.maximum(1.0, np.abs(nv_).max(axis=1))
    eq = np.abs(gq - nq).max(axis=1) / np.maximum(1.0, np.abs(nq).max(axis=1))
    for e in range(n):
        records.append((ev[e], eq[e], oncon[e], t, e))
    e = int(np.argmax(ev))
    if ev[e] > worst[0]:
        worst = (float(ev[e]), dict(t=t, e=e, q=q.copy(), v=v.copy(), p=p.copy(), w=w.copy(), a=a.copy()))
rec = np.array([(r[0], r[1], r[2]) for r in records])
print(f"{args.task}: {len(rec)} single steps; qvel rel err quantiles 50/90/99/99.9/max:",
      " ".join(f"{x:.2e}" for x in np.quantile(rec[:, 0], [0.5, 0.9, 0.99, 0.999, 1.0])))
print("   qpos rel err quantiles:", " ".join(f"{x:.2e}" for x in np.quantile(rec[:, 1], [0.5, 0.9, 0.99, 0.999, 1.0])))
print("   fraction with qvel err > 1e-4:", float((rec[:, 0] > 1e-4).mean()), " mean ncon", rec[:, 2].mean(), "max ncon", rec[:, 2].max())
for lo, hi in [(0, 1), (1, 4), (4, 8), (8, 16), (16, 100)]:
    sel = (rec[:, 2] >= lo) & (rec[:, 2] < hi)
    if sel.any():
        print(f"   ncon in [{lo},{hi}): count {int(sel.sum())}, median err {np.median(rec[sel, 0]):.2e}, max {rec[sel, 0].max():.2e}, "
              f">1e-4: {float((rec[sel, 0] > 1e-4).mean()):.3f}")

print('forward qacc rel err > 1e-3:', nbig, 'of', ntot, dict(cls))
# dig into the worst state
W = worst[1]
print(f"worst: err {worst[0]:.3e} at t={W['t']} env={W['e']}")
e = W["e"]
ora.set_state(W["q"], W["v"], W["p"]); ora.set_warmstart(W["w"])
env._set_raw_state(W["q"], W["v"], W["p"]); env.set_warmstart(W["w"])
ora.set_state(W["q"], W["v"], W["p"])
gq_, gncon, gnefc, _ = env.solver()
oq_, oncon, onefc, _ = ora.solver()
print("  ncon gpu/oracle", gncon[e], oncon[e], " nefc", gnefc[e], onefc[e])
oc = ora.contacts(e)
gc = env.debug_scratch(L["G_CON"], 16 * L["MAXCON"])[e].reshape(-1, 16)[: gncon[e]]
np.set_printoptions(precision=6, suppress=True, linewidth=200)
gi = np.zeros((n, L["GI_TOTAL"]), dtype=np.int32)
env.abi.lib.adroit_debug_iscratch(env._h, gi.ctypes.data, None)
print("  newton iters gpu", gi[e, 2], " noslip sweeps", gi[e, 3])
gname = lambda g: model.names["geom"][g] if hasattr(model, "names") and "geom" in model.names else str(g)
for k in range(max(len(oc), len(gc))):
    o = oc[k] if k < len(oc) else None
    g = gc[k] if k < len(gc) else None
    gg = gi[e, L["GI_CON"] + 4 * k: L["GI_CON"] + 4 * k + 3] if k < len(gc) else None
    print("   contact", k, "geoms", gg, "types", None if gg is None else (model.A["geom_type"][gg[0]], model.A["geom_type"][gg[1]]))
    if o is not None:
        print("      oracle dist %.3e pos %s n %s" % (o[0], o[1:4], o[4:7]))
    if g is not None:
        print("      gpu    dist %.3e pos %s n %s" % (g[0], g[1:4], g[4:7]))
nv = env.nv
qas = env.debug_scratch(L["G_DBG_QAS"], nv)[e]
d = ora.dynamics(e)
print("  qacc_smooth rel err", np.abs(qas - d["qacc_smooth"]).max() / max(1, np.abs(d["qacc_smooth"]).max()))
print("  qacc   gpu", gq_[e])
print("  qacc   ora", oq_[e])
print("  qacc rel err", np.abs(gq_[e] - oq_[e]).max() / max(1, np.abs(oq_[e]).max()))
qfc = env.debug_scratch(L["G_DBG_QFC"], nv)[e]
print("  qfrc_constraint err", np.abs(qfc - d["qfrc_constraint"]).max(), " scale", np.abs(d["qfrc_constraint"]).max())
env.close()
