"""Replays tests/golden/shim_*.npz -- outputs of the reference's OWN task code (hammer_v0.py / door_v0.py / pen_v0.py /
relocate_v0.py run unmodified on the shim of oracle/ref_shim.py, see tools/make_shim_fixtures.py) -- through the CUDA
path (AdroitVecEnv, C ABI).  The reference tree is not on the GPU box; the fixtures are its stand-in."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CONFIGS = [("hammer-v0", None), ("hammer-v0", "mass"), ("hammer-v0", "pos"), ("hammer-v0", "size"), ("door-v0", None),
           ("pen-v0", None), ("relocate-v0", None)]
TASK = {"hammer-v0": "hammer", "door-v0": "door", "pen-v0": "pen", "relocate-v0": "relocate"}


def _load(env_id, var):
    return dict(np.load(os.path.join(GOLD, f"shim_{TASK[env_id]}" + (f"_{var}" if var else "") + ".npz")))


@pytest.mark.parametrize("env_id,var", CONFIGS)
def test_cuda_step_reproduces_reference_task_code_outputs(env_id, var):
    """fp64 build, free-running from the reference's reset state under the reference's action sequence: observation,
    reward, goal_achieved and done of every step; then get_env_state (reference keys, lagged frames) and
    set_env_state + get_obs."""
    import torch
    from mj_envs_b200.vec_env import AdroitVecEnv
    fx = _load(env_id, var)
    E, T = fx["action"].shape[:2]
    env = AdroitVecEnv(env_id, E, seed=int(fx["seed"]), fp64=True, variation_type=var)
    env.reset()                                                                    # zero warm start, like sim.reset()
    env._set_raw_state(np.zeros((E, env.nv)), np.zeros((E, env.nv)), fx["par"])
    assert np.abs(env.get_obs().cpu().numpy() - fx["reset_obs"]).max() < 1e-6
    for t in range(T):
        obs, rew, done, info = env.step(torch.from_numpy(fx["action"][:, t]))
        q, v, _ = env._raw_state()
        assert np.abs(q - fx["qpos"][:, t]).max() < 1e-7 and np.abs(v - fx["qvel"][:, t]).max() < 1e-5, t     # fp64 free run stays on the oracle's trajectory
        assert np.abs(obs.cpu().numpy() - fx["obs"][:, t]).max() < 1e-5, t
        assert np.abs(rew.cpu().numpy() - fx["reward"][:, t]).max() < 1e-5 * max(1.0, np.abs(fx["reward"][:, t]).max()), t
        assert np.array_equal(info["goal_achieved"].cpu().numpy(), fx["goal"][:, t]), t
        assert np.array_equal(done.cpu().numpy(), fx["done"][:, t]), t
    sd = env.get_env_state()
    keys = sorted(k[3:] for k in fx if k.startswith("sd_") and k != "sd_obs")
    assert sorted(sd) == keys
    for k in keys:
        assert isinstance(sd[k], torch.Tensor) and sd[k].shape[0] == E
        assert np.abs(sd[k].numpy() - fx["sd_" + k]).max() < 1e-6, k                # incl. the one-substep-lagged site / body frames
    obs_before = env.get_obs()
    sd2 = env.get_env_state()                                                       # no side effect: a second read and get_obs are unchanged
    assert all(torch.equal(sd[k], sd2[k]) for k in sd) and torch.equal(obs_before, env.get_obs())
    env.set_env_state({k: fx["sd_" + k] for k in keys})
    assert np.abs(env.get_obs().cpu().numpy() - fx["sd_obs"]).max() < 1e-5
    assert env.overflow().sum() == 0
    env.close()


@pytest.mark.parametrize("env_id", ["hammer-v0", "door-v0", "pen-v0", "relocate-v0"])
@pytest.mark.parametrize("fp64", [True, False])
def test_cuda_path_on_crafted_states_reaches_the_reference_bonus_goal_done_branches(env_id, fp64):
    import torch
    from mj_envs_b200.vec_env import AdroitVecEnv
    fx = _load(env_id, None)
    keys = sorted(k[len("crafted_sd_"):] for k in fx if k.startswith("crafted_sd_"))
    n = fx["crafted_reward"].shape[0]
    env = AdroitVecEnv(env_id, n, fp64=fp64)
    env.reset()
    env.set_env_state({k: fx["crafted_sd_" + k] for k in keys})
    tol = 1e-5 if fp64 else 2e-4
    assert np.abs(env.get_obs().cpu().numpy() - fx["crafted_obs0"]).max() < tol
    obs, rew, done, info = env.step(torch.zeros((n, env.nu)))
    assert np.array_equal(info["goal_achieved"].cpu().numpy(), fx["crafted_goal"]) and info["goal_achieved"].any()
    assert np.array_equal(done.cpu().numpy(), fx["crafted_done"])
    assert np.abs(rew.cpu().numpy() - fx["crafted_reward"]).max() < (1e-4 if fp64 else 2e-3) * max(1.0, np.abs(fx["crafted_reward"]).max())
    assert np.abs(obs.cpu().numpy() - fx["crafted_obs"]).max() < (1e-5 if fp64 else 5e-3)
    env.close()


@pytest.mark.parametrize("env_id,var", CONFIGS)
def test_reset_draws_follow_the_reference_order_and_ranges(env_id, var):
    """num_envs = 1, four resets = episodes 0..3 of global env 0: the parameters equal what the reference's reset_model wrote
    when its np_random.uniform(low, high) calls were fed the same Philox uniforms (order + ranges), and the reset observation
    is the reference's."""
    from mj_envs_b200.vec_env import AdroitVecEnv
    fx = _load(env_id, var)
    env = AdroitVecEnv(env_id, 1, seed=int(fx["seed"]), fp64=True, variation_type=var)
    for ep in range(fx["par"].shape[0]):
        obs, info = env.reset()
        assert info == {} and obs.shape == (1, env.obs_dim)
        assert np.abs(env._raw_state()[2][0] - fx["par"][ep]).max() < 1e-12
        assert np.abs(obs.cpu().numpy()[0] - fx["reset_obs"][ep]).max() < 1e-6
    env.close()
