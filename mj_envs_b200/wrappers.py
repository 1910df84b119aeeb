"""Batched counterparts of the reference's gym adaptation layer:
mj_envs_vision/utils/wrappers.py (CustomPixelObservationWrapper, StateActionSpec), utils/helpers.py (make_env / step /
reset) and the `paths` structure consumed by `evaluate_success` (hammer_v0.py:167-175).  Everything is batched on dim 0
and stays on the device; nothing here touches the oracle.
"""
from __future__ import annotations

import numpy as np
import torch

from .vec_env import AdroitVecEnv, ENV_IDS

PIXELS_KEY = "pixels"   # wrappers.py:9
STATE_KEY = "state"     # wrappers.py:10


class StateActionSpec:   # wrappers.py:13-16
    def __init__(self, action_space, observation_space):
        self.action_dim = action_space.shape[0]
        self.observation_dim = observation_space.shape[0]


def images_to_observation(images: torch.Tensor, bit_depth: int = 5) -> torch.Tensor:
    """[EXT] PlaNet `_images_to_observation` as used by the reference (utils/train.py:122,189; algos/baselines.py:313),
    batched: u8/float [N, 64, 64, 3] in [0, 255] -> float [N, 3, 64, 64] quantised to `bit_depth` bits, centred in
    [-0.5, 0.5).  (The reference resizes its 128x128 crop to 64x64 first; the render kernel already delivers 64x64.  The
    dequantisation noise PlaNet adds at training time is the caller's business.)"""
    x = images.to(torch.float32).permute(0, 3, 1, 2)
    return torch.floor(x / 2 ** (8 - bit_depth)) / 2 ** bit_depth - 0.5


class BatchedPixelObservationWrapper:
    """CustomPixelObservationWrapper (wrappers.py:32-76) over an AdroitVecEnv: keeps a dict observation {state, pixels},
    returns the entry `obs_key` as a float tensor, speaks the new gym step API the reference converts to with
    StepAPICompatibility(output_truncation_bool=True) (wrappers.py:34): (obs, reward, terminated, truncated, info), and
    repeats an action `action_repeat` times accumulating the reward (wrappers.py:57-70; the reference's own loop would
    crash on `items[1] += ...` for action_repeat > 1 -- SURVEY.md App. A.8 Q7 -- the accumulation it intends is what is
    implemented, per env, stopping at the env's first done).  Pixels are only rendered when they are asked for."""

    def __init__(self, env: AdroitVecEnv, obs_key: str = PIXELS_KEY, action_repeat: int = 1, max_episode_length: int | None = None):
        if obs_key not in (PIXELS_KEY, STATE_KEY):
            raise Exception(f"Unsupported state type '{obs_key}'")     # helpers.py:77
        self.env = env
        self.unwrapped = env
        self.obs_key = obs_key
        self.action_repeat = int(action_repeat)
        self.max_episode_length = max_episode_length or env.max_episode_steps   # the reference hard-codes 200 (wrappers.py:38)
        self.env_spec = StateActionSpec(env.action_space, env.observation_space)
        self.action_space = env.action_space
        self.timer = 0
        self.curr_obs = {}
        self.reset()
        shape = tuple(self.curr_obs[obs_key].shape[1:])
        from .vec_env import Box
        self.observation_space = Box(0, 255 if obs_key == PIXELS_KEY else 1, shape)

    def _observe(self, state):
        self.curr_obs = {STATE_KEY: state}
        if self.obs_key == PIXELS_KEY:
            self.curr_obs[PIXELS_KEY] = self.env.render().to(torch.float32)   # HeadlessObserver.render returns float in [0, 255]

    def reset(self):
        self.timer = 0
        state, info = self.env.reset()
        self._observe(state)
        return self.curr_obs[self.obs_key].clone(), info

    def step(self, action):
        state, reward, done, info = self.env.step(action)
        self.timer += 1
        alive = ~done
        for _ in range(self.action_repeat - 1):
            if self.timer > self.max_episode_length:
                break
            s2, r2, d2, i2 = self.env.step(action)
            self.timer += 1
            reward = reward + torch.where(alive, r2, torch.zeros_like(r2))      # accumulate rewards (wrappers.py:68)
            state = torch.where(alive[:, None], s2, state)                      # update obs (wrappers.py:67)
            info = {"goal_achieved": info["goal_achieved"] | (i2["goal_achieved"] & alive)}
            done = done | d2
            alive = alive & ~d2
        self._observe(state)
        terminated = done
        truncated = torch.zeros_like(done)       # TimeLimit is stripped by make_env (helpers.py:68-69)
        return self.curr_obs[self.obs_key].clone(), reward, terminated, truncated, info

    def get_pixels(self) -> torch.Tensor:    # wrappers.py:72-73
        if PIXELS_KEY not in self.curr_obs:
            self.curr_obs[PIXELS_KEY] = self.env.render().to(torch.float32)
        return self.curr_obs[PIXELS_KEY].clone()

    def get_state(self) -> torch.Tensor:     # wrappers.py:75-76
        return self.curr_obs[STATE_KEY].clone()

    def close(self):
        self.env.close()


# ---------------------------------------------------------------- utils/helpers.py:41-78
def is_from_adroit_suite(env_name: str) -> bool:
    return env_name in ENV_IDS


def make_env(config, num_envs: int = 1, device=0):
    """helpers.make_env (helpers.py:56-78) for the batched simulator.  Reads the same config keys: env_name, nogui,
    state_type in {"observation" (pixels), "vector" (state)}, variation_type, seed, action_repeat."""
    assert is_from_adroit_suite(config.env_name)
    env = AdroitVecEnv(config.env_name, num_envs, device=device, seed=int(getattr(config, "seed", 0) or 0),
                       variation_type=getattr(config, "variation_type", None))
    st = getattr(config, "state_type", "vector")
    if st == "observation":
        return BatchedPixelObservationWrapper(env)                      # make_env never forwards action_repeat (Q7)
    if st == "vector":
        return BatchedPixelObservationWrapper(env, obs_key=STATE_KEY)
    raise Exception(f"Unsupported state type '{st}'")


def reset(env):   # helpers.py:41-42
    return env.reset()


def step(env, action: torch.Tensor):
    """helpers.step (helpers.py:44-54): (obs, reward, done, success) with success = info['goal_achieved']."""
    res = env.step(action)
    obs, reward, done = res[:3]
    return obs, reward, done, res[-1]["goal_achieved"]


# ---------------------------------------------------------------- paths (hammer_v0.py:167-175 and twins; eval.py:69-97)
from .vec_env import evaluate_success  # noqa: E402,F401  (hammer_v0.py:167-175 and twins; one implementation, re-exported here)


def collect_paths(env: AdroitVecEnv, policy, horizon: int | None = None, record_state: bool = False):
    """One episode of every env of the batch under `policy(obs) -> action`, returned as the reference's `paths`: a list of
    N dicts {observations [T, obs_dim], actions [T, nu], rewards [T], env_infos: {goal_achieved [T]}} (+ `init_state_dict`
    = get_env_state() of the batch entry when record_state) -- what evaluate_success and mjrl-style tooling consume."""
    horizon = horizon or env.max_episode_steps
    obs, _ = env.reset()
    init = env.get_env_state() if record_state else None
    O, A, R, G = [], [], [], []
    for _ in range(horizon):
        a = policy(obs)
        O.append(obs)
        obs, r, _, info = env.step(a)
        A.append(torch.as_tensor(a, device=obs.device, dtype=torch.float32))
        R.append(r)
        G.append(info["goal_achieved"])
    O, A, R, G = (torch.stack(x, dim=1).cpu().numpy() for x in (O, A, R, G))
    paths = []
    for e in range(env.num_envs):
        p = dict(observations=O[e], actions=A[e], rewards=R[e], env_infos=dict(goal_achieved=G[e]))
        if init is not None:
            p["init_state_dict"] = {k: np.asarray(v)[e].copy() for k, v in init.items()}
        paths.append(p)
    return paths
