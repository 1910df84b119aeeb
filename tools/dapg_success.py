"""Success rates of the reference's pretrained DAPG policies (trained on real MuJoCo; weights extracted to tests/golden/dapg_*.npz)
on the fp32 CUDA path, the fp64 CUDA build and the fp64 CPU oracle -- the one external pin of the PHYSICS this box allows.
evaluate_success semantics (hammer_v0.py:167-175): an episode succeeds if goal_achieved holds on more than 25 (pen: 20) steps.
Usage (GPU box): python tools/dapg_success.py > profiles/dapg_success.json"""
import json, os, sys
import numpy as np
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import torch  # noqa: E402
from mj_envs_b200.vec_env import AdroitVecEnv  # noqa: E402
from mj_envs_b200.mjcf import load_compiled  # noqa: E402
from oracle.oracle import OracleSim  # noqa: E402
GOLD = os.path.join(R, "tests", "golden")


def weights(task):
    return {k: np.asarray(v, dtype=np.float64) for k, v in np.load(os.path.join(GOLD, f"dapg_{task}.npz")).items() if np.asarray(v).dtype.kind == "f"}


def act_np(z, obs):
    h = (obs - z["in_shift"]) / (z["in_scale"] + 1e-8)
    h = np.tanh(h @ z["W0"].T + z["b0"]); h = np.tanh(h @ z["W1"].T + z["b1"])
    return (h @ z["W2"].T + z["b2"]) * z["out_scale"] + z["out_shift"]


out = {"protocol": "one full episode per env from reset (seed 3), deterministic policy mean, success = sum(goal_achieved) > 25 (pen: 20)", "rows": []}
for task, fs in (("hammer", None), ("door", 5), ("door", 1), ("pen", None), ("relocate", None)):
    z = weights(task)
    thr = 20 if task == "pen" else 25
    row = {"task": f"{task}-v0", "frame_skip": fs if fs else 5}
    for name, n, fp64 in (("cuda_fp32", 1024, False), ("cuda_fp64", 256, True)):
        env = AdroitVecEnv(f"{task}-v0", n, seed=3, frame_skip=fs, fp64=fp64)
        zt = {k: torch.tensor(v, dtype=torch.float32, device=env.device) for k, v in z.items()}
        obs, _ = env.reset()
        goals = torch.zeros(n, device=env.device); ret = torch.zeros(n, device=env.device)
        for _ in range(env.horizon):
            h = (obs - zt["in_shift"]) / (zt["in_scale"] + 1e-8)
            h = torch.tanh(h @ zt["W0"].T + zt["b0"]); h = torch.tanh(h @ zt["W1"].T + zt["b1"])
            obs, r, d, info = env.step((h @ zt["W2"].T + zt["b2"]) * zt["out_scale"] + zt["out_shift"])
            goals += info["goal_achieved"]; ret += r
        row[name] = {"envs": n, "success": float((goals > thr).float().mean()), "mean_return": float(ret.mean()), "overflow": int(env.overflow().sum())}
        env.close()
    m = load_compiled(task)
    if fs:
        m.A["sizes"] = m.A["sizes"].copy(); m.A["sizes"][9] = fs
    n = 64
    s = OracleSim(m, n, seed=3)
    obs = s.reset(); goals = np.zeros(n); ret = np.zeros(n)
    for _ in range(s.horizon):
        obs, r, d, g = s.step(act_np(z, obs.astype(np.float64)).astype(np.float32))
        goals += g; ret += r
    row["oracle_fp64"] = {"envs": n, "success": float((goals > thr).mean()), "mean_return": float(ret.mean())}
    out["rows"].append(row)
    print(row, file=sys.stderr)
print(json.dumps(out, indent=1))
