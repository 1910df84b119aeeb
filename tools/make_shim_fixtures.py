"""Generates tests/golden/shim_*.npz: outputs of the reference's OWN task code
(/root/reference/mj_envs_vision/hand_manipulation_suite/{hammer,door,pen,relocate}_v0.py, imported unmodified on
the stub mujoco_py / mjrl / gym of oracle/ref_shim.py, physics = the fp64 oracle) on seeded action sequences:
4 episodes x 50 steps for every task and for the three hammer variations, plus crafted states that reach the
bonus / goal / done branches.  The reference tree does not exist on the GPU box, so the vectors are committed;
tests/test_reference_shim.py re-derives them live whenever /root/reference is present and checks the oracle's
restated task layer (oracle/oracle_task.inc) against them; tests/test_gpu_shim_replay.py replays them through the
CUDA path (AdroitVecEnv).

Usage: python tools/make_shim_fixtures.py"""
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
from oracle import ref_shim as R  # noqa: E402

EPISODES, STEPS, SEED = 4, 50, 3


def name(env_id, var):
    return f"shim_{R.TASK[env_id]}" + (f"_{var}" if var else "")


def generate(env_id, var):
    fix, rep = R.rollout_against_restatement(env_id, var, seed=SEED, episodes=EPISODES, steps=STEPS)
    if var is None:
        cf, crep = R.crafted_against_restatement(env_id)
        fix.update(cf)
        rep.update({"crafted_" + k: v for k, v in crep.items()})
    return fix, rep


if __name__ == "__main__":
    out = os.path.join(ROOT, "tests", "golden")
    for (env_id, var) in R.N_DRAWS:
        fix, rep = generate(env_id, var)
        np.savez_compressed(os.path.join(out, name(env_id, var) + ".npz"), **fix)
        print(name(env_id, var), {k: (float(f"{v:.3g}") if isinstance(v, float) else v) for k, v in rep.items()})
    print("registered by the reference:", R.REGISTERED)
