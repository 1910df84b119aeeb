"""GPU tests at the behavioural / statistical level and at BASELINE.json's full sizes (size-independent properties)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _policy(task, device):
    import torch
    z = {k: torch.tensor(np.asarray(v), dtype=torch.float32, device=device) for k, v in np.load(os.path.join(GOLD, f"dapg_{task}.npz")).items()
         if np.asarray(v).dtype.kind == "f"}

    def act(obs):
        h = (obs - z["in_shift"]) / (z["in_scale"] + 1e-8)
        h = torch.tanh(h @ z["W0"].T + z["b0"])
        h = torch.tanh(h @ z["W1"].T + z["b1"])
        return (h @ z["W2"].T + z["b2"]) * z["out_scale"] + z["out_shift"]
    return act


@pytest.mark.parametrize("task", ["door", "hammer", "pen", "relocate"])
def test_dapg_policies_trained_on_mujoco_solve_their_tasks_on_the_cuda_path(task):
    """Known-answer test from the reference's own artefacts (algos/dapg_pretrained/*.pickle -> tests/golden): policies
    trained on real MuJoCo physics must succeed on the fp32 CUDA path (door at the upstream frame_skip=5 it was trained at)."""
    import torch
    from mj_envs_b200.vec_env import AdroitVecEnv
    n = 256
    env = AdroitVecEnv(f"{task}-v0", n, seed=3, frame_skip=5 if task == "door" else None)
    act = _policy(task, env.device)
    obs, _ = env.reset()
    goals = torch.zeros(n, device=env.device)
    ret = torch.zeros(n, device=env.device)
    for _ in range(env.horizon):
        obs, r, d, info = env.step(act(obs))
        goals += info["goal_achieved"]
        ret += r
    thr = 20 if task == "pen" else 25
    success = (goals > thr).float().mean().item()
    assert env.overflow().sum() == 0
    assert success >= 0.75, (success, ret.mean().item())
    paths = [{"env_infos": {"goal_achieved": np.ones(thr + 1)}}, {"env_infos": {"goal_achieved": np.zeros(50)}}]
    assert env.evaluate_success(paths) == 50.0          # hammer_v0.py:167-175 semantics
    env.close()


@pytest.mark.parametrize("task,frame_skip", [("door", 1), ("relocate", 5)])
def test_random_policy_returns_are_statistically_indistinguishable_from_the_oracle(task, frame_skip):
    """Free-running fp32 CUDA rollouts vs fp64 oracle rollouts under the same random action stream: two-sample KS on the
    episode returns (independent env seeds on the two sides would add sampling noise; here only physics differs)."""
    import torch
    from scipy import stats
    from mj_envs_b200.vec_env import AdroitVecEnv
    from oracle.oracle import OracleSim
    n, T = 256, 60
    env = AdroitVecEnv(f"{task}-v0", n, seed=21)
    ora = OracleSim(task, n, seed=21)
    env.reset(); ora.reset()
    rng = np.random.default_rng(5)
    gret, oret = np.zeros(n), np.zeros(n)
    for t in range(T):
        a = rng.uniform(-1, 1, (n, env.nu)).astype(np.float32)
        _, r, _, _ = env.step(torch.from_numpy(a))
        _, orw, _, _ = ora.step(a)
        gret += r.cpu().numpy(); oret += orw
    ks = stats.ks_2samp(gret, oret)
    assert ks.pvalue > 0.05, ks
    assert abs(gret.mean() - oret.mean()) < 0.05 * abs(oret.mean()) + 0.5
    env.close()


@pytest.mark.parametrize("env_id,n", [("door-v0", 4096), ("pen-v0", 16384), ("relocate-v0", 8192), ("hammer-v0", 8192)])
def test_full_size_configs_are_deterministic_finite_and_within_capacity(env_id, n):
    """BASELINE.json configs 2-5 at full per-GPU size: same seed + same actions => bitwise identical results; no NaN;
    no contact/row capacity overflow; rewards inside the analytic bounds of the reward functions; resets restore qpos0."""
    import torch
    from mj_envs_b200.vec_env import AdroitVecEnv
    outs = []
    for rep in range(2):
        env = AdroitVecEnv(env_id, n, seed=77, auto_reset=True)
        g = torch.Generator(device=env.device); g.manual_seed(1)
        obs, _ = env.reset()
        tot = torch.zeros(n, device=env.device)
        for t in range(6):
            a = torch.rand((n, env.nu), device=env.device, generator=g) * 2 - 1
            obs, r, d, info = env.step(a)
            tot += r
        assert torch.isfinite(obs).all() and torch.isfinite(tot).all()
        assert env.overflow().sum() == 0
        assert r.max().item() < 120 and r.min().item() > -60
        outs.append((obs.clone(), tot.clone()))
        env.close()
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


def test_reset_mask_only_touches_selected_envs():
    import torch
    from mj_envs_b200.vec_env import AdroitVecEnv
    n = 64
    env = AdroitVecEnv("hammer-v0", n, seed=5)
    env.reset()
    for _ in range(3):
        env.step(torch.rand((n, env.nu)) * 2 - 1)
    q0, v0, p0 = env._raw_state()
    mask = torch.zeros(n, dtype=torch.uint8); mask[::2] = 1
    env.reset(mask)
    q1, v1, p1 = env._raw_state()
    assert np.array_equal(q1[1::2], q0[1::2]) and np.array_equal(v1[1::2], v0[1::2])
    assert np.abs(q1[::2]).max() == 0 and np.abs(v1[::2]).max() == 0
    env.close()
