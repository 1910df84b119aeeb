"""Gym-adaptation layer: batched counterparts of utils/wrappers.py, utils/helpers.py and the
`paths` structure of evaluate_success."""
import types

import numpy as np
import pytest
import torch

from mj_envs_b200 import wrappers as W


def test_images_to_observation_is_planets_preprocessing():
    img = torch.arange(0, 256, dtype=torch.uint8).repeat(48)[: 64 * 64 * 3].reshape(1, 64, 64, 3)
    out = W.images_to_observation(img, bit_depth=5)
    assert out.shape == (1, 3, 64, 64)
    ref = np.floor(img.numpy().astype(np.float64) / 8) / 32 - 0.5       # [EXT PlaNet env.py] preprocess_observation_
    assert np.allclose(out.numpy(), ref.transpose(0, 3, 1, 2))
    assert out.min() >= -0.5 and out.max() < 0.5


def test_evaluate_success_thresholds_match_the_reference():
    mk = lambda n: dict(env_infos=dict(goal_achieved=np.array([True] * n + [False] * (200 - n))))
    paths = [mk(26), mk(25), mk(21), mk(0)]
    assert W.evaluate_success("hammer-v0", paths) == 25.0     # > 25 (hammer_v0.py:172)
    assert W.evaluate_success("pen-v0", paths) == 75.0        # > 20 (pen_v0.py:185)


def test_make_env_rejects_unknown_state_type_without_touching_the_gpu():
    cfg = types.SimpleNamespace(env_name="HalfCheetah-v2", state_type="vector")
    with pytest.raises(AssertionError):                       # helpers.py:57
        W.make_env(cfg)


@pytest.mark.gpu
def test_wrapper_speaks_the_new_step_api_and_returns_the_requested_key():
    cfg = types.SimpleNamespace(env_name="relocate-v0", state_type="observation", variation_type=None, seed=3, nogui=True)
    env = W.make_env(cfg, num_envs=8)
    obs, info = W.reset(env)
    assert obs.shape == (8, 64, 64, 3) and obs.dtype == torch.float32 and info == {}
    a = torch.zeros((8, env.env_spec.action_dim))
    out = env.step(a)
    assert len(out) == 5 and out[0].shape == (8, 64, 64, 3) and out[1].shape == (8,) and out[2].dtype == torch.bool
    assert env.get_state().shape == (8, env.env_spec.observation_dim) and env.get_pixels().shape == (8, 64, 64, 3)
    o, r, d, s = W.step(env, a)                                # helpers.step: (obs, reward, done, success)
    assert s.shape == (8,) and s.dtype == torch.bool
    cfg.state_type = "vector"
    env2 = W.make_env(cfg, num_envs=8)
    assert W.reset(env2)[0].shape == (8, 39)
    cfg.state_type = "bogus"
    with pytest.raises(Exception, match="Unsupported state type"):
        W.make_env(cfg, num_envs=2)
    env.close(); env2.close()


@pytest.mark.gpu
def test_action_repeat_accumulates_rewards():
    from mj_envs_b200.vec_env import AdroitVecEnv
    a = torch.rand((16, 24)) * 2 - 1
    w = W.BatchedPixelObservationWrapper(AdroitVecEnv("pen-v0", 16, seed=5), obs_key=W.STATE_KEY, action_repeat=3)
    e = AdroitVecEnv("pen-v0", 16, seed=5)
    w.reset(); e.reset(); e.reset()     # the wrapper's constructor already reset once (wrappers.py:43 does the same)
    obs, rew, term, trunc, info = w.step(a)
    tot = torch.zeros(16, device=rew.device)
    alive = torch.ones(16, dtype=torch.bool, device=rew.device)
    for k in range(3):
        o, r, d, i = e.step(a)
        tot += torch.where(alive, r, torch.zeros_like(r)) if k else r
        alive = alive & ~d if k else ~d
    assert torch.allclose(rew, tot) and not trunc.any()
    w.close(); e.close()


@pytest.mark.gpu
def test_collected_paths_agree_with_the_device_side_episode_statistics():
    from mj_envs_b200.vec_env import AdroitVecEnv
    env = AdroitVecEnv("door-v0", 32, seed=2)
    g = torch.Generator(device=env.device); g.manual_seed(0)
    paths = W.collect_paths(env, lambda obs: torch.rand((32, env.nu), device=env.device, generator=g) * 2 - 1, horizon=60, record_state=True)
    assert len(paths) == 32 and paths[0]["observations"].shape == (60, 39) and paths[0]["actions"].shape == (60, 28)
    assert set(paths[0]["init_state_dict"]) == {"qpos", "qvel", "door_body_pos"}     # door_v0.py:128
    stats = env.reduce_episode(flush_running=True, all_reduce=False)
    assert stats[3] == 32
    assert abs(stats[0] - sum(p["rewards"].sum() for p in paths)) < 1e-2 * max(1.0, abs(stats[0]))
    assert W.evaluate_success("door-v0", paths) == env.evaluate_success(paths) == 100.0 * stats[2] / 32
    env.close()
