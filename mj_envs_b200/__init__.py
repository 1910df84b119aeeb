"""mj_envs_b200 — B200-native batched simulator for the Adroit hand-manipulation tasks (drop-in for the step path of
bilkitty/mj_envs: door-v0, hammer-v0, pen-v0, relocate-v0)."""
from .mjcf import Model, load_compiled, load_mjcf  # noqa: F401


def __getattr__(name):  # lazy: importing the package must not require torch / CUDA
    if name in ("AdroitVecEnv", "make", "ENV_IDS", "MAX_EPISODE_STEPS"):
        from . import vec_env
        return getattr(vec_env, name)
    raise AttributeError(name)
