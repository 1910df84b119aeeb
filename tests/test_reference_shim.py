"""Pins the task layer to the reference's OWN code.

`oracle/ref_shim.py` runs /root/reference/mj_envs_vision/hand_manipulation_suite/{hammer,door,pen,relocate}_v0.py
unmodified over stub mujoco_py / mjrl / gym modules whose MjSim is the fp64 oracle.  Live tests (skipped where the
reference tree is absent, e.g. the GPU box) compare the reference code's step / get_obs / reset_model /
get_env_state / set_env_state / evaluate_success with the oracle's restated task layer (oracle/oracle_task.inc) and
with the committed fixtures; fixture tests (run anywhere) replay tests/golden/shim_*.npz through the oracle.
Tolerances: observations are float32 on both sides (bit-equal up to a sign of zero); rewards 1e-12 relative in fp64."""
import os

import numpy as np
import pytest

from oracle import ref_shim as R
from oracle.oracle import OracleSim
from mj_envs_b200.mjcf import load_compiled

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CONFIGS = list(R.N_DRAWS)
live = pytest.mark.skipif(not R.available(), reason="/root/reference is not present on this box")
# reset_model draws, in the reference's order (hammer_v0.py:109-125, door_v0.py:107-109,
# pen_v0.py:120-121, relocate_v0.py:89-93)
RANGES = {("hammer-v0", None): [(0.1, 0.25)], ("hammer-v0", "mass"): [(0.1, 0.25), (0.05, 2.5)],
          ("hammer-v0", "pos"): [(0.1, 0.25), (-0.24, -0.10)], ("hammer-v0", "size"): [(0.1, 0.25), (0.01, 0.04), (0.02, 0.08)],
          ("door-v0", None): [(-0.3, -0.2), (0.25, 0.35), (0.252, 0.35)], ("pen-v0", None): [(-1, 1), (-1, 1)],
          ("relocate-v0", None): [(-0.15, 0.15), (-0.15, 0.3), (-0.2, 0.2), (-0.2, 0.2), (0.15, 0.35)]}


def _name(env_id, var):
    return f"shim_{R.TASK[env_id]}" + (f"_{var}" if var else "")


def _load(env_id, var):
    return dict(np.load(os.path.join(GOLD, _name(env_id, var) + ".npz")))


def _assert_report(rep):
    assert rep["par"] <= 1e-15          # Philox uniforms through the reference's np_random.uniform calls = the oracle's own draws
    assert rep["state"] == 0.0          # same physics on both sides: any later deviation is the task layer's
    assert rep["reset_obs"] <= 1e-7 and rep["obs"] <= 1e-7 and rep["sd_obs"] <= 1e-7
    assert rep["obs_exact"] >= 0.99 * rep["obs_total"]
    assert rep["reward"] <= 1e-12
    assert rep["flags"] == 0            # goal_achieved and done, every step
    assert rep["sd"] <= 1e-15           # get_env_state values (keys are asserted inside the rollout)


@live
@pytest.mark.parametrize("env_id,var", CONFIGS)
def test_reference_task_code_equals_restated_task_layer(env_id, var):
    fix, rep = R.rollout_against_restatement(env_id, var, seed=11, episodes=2, steps=30)
    _assert_report(rep)
    assert [tuple(x) for x in fix["draws"][0]] == [tuple(map(float, r)) for r in RANGES[(env_id, var)]]
    # action space and registration, as the reference's own modules report them
    assert np.all(fix["action_low"] == -1) and np.all(fix["action_high"] == 1)                    # hammer_v0.py:51-52
    from mj_envs_b200.vec_env import ENV_IDS, MAX_EPISODE_STEPS
    assert set(R.REGISTERED) == set(ENV_IDS)
    assert {k: v["max_episode_steps"] for k, v in R.REGISTERED.items()} == MAX_EPISODE_STEPS       # mj_envs_vision/__init__.py:4-29


@live
@pytest.mark.parametrize("env_id", list(R.ENV_CLASSES))
def test_crafted_states_reach_bonus_goal_and_done_branches(env_id):
    fix, rep = R.crafted_against_restatement(env_id)
    assert rep["obs"] <= 1e-7 and rep["reward"] <= 1e-12 and rep["flags"] == 0
    assert rep["goals"] >= 1
    assert rep["dones"] == (1 if env_id == "pen-v0" else 0)
    gold = _load(env_id, None)
    for k, v in fix.items():                  # and the committed fixture is what the reference produces today
        assert np.array_equal(v, gold[k]), k


@live
def test_committed_fixture_regenerates_bit_identically():
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_shim_fixtures", os.path.join(GOLD, "..", "..", "tools", "make_shim_fixtures.py"))
    gen = importlib.util.module_from_spec(spec); spec.loader.exec_module(gen)
    fix, rep = R.rollout_against_restatement("door-v0", None, seed=gen.SEED, episodes=gen.EPISODES, steps=gen.STEPS)
    gold = _load("door-v0", None)
    for k, v in fix.items():
        assert np.array_equal(v, gold[k]), k


@live
def test_reference_constructor_quirks_are_reproduced():
    """App. A.8 Q4 / Q5: MujocoEnv.__init__ steps once with unscaled zero ctrl; door / pen / relocate call reset_model inside
    __init__ with placeholder ids (draws land on the world body / site 0) -- the shim lets them and the next reset erases them."""
    for env_id, ctor_draws in (("hammer-v0", 0), ("door-v0", 3), ("pen-v0", 2), ("relocate-v0", 5)):
        env = R.make_reference_env(env_id)
        assert len(env.np_random.calls) == ctor_draws
        assert env.sim.nsteps == env.frame_skip                      # the one step(zeros) of MujocoEnv.__init__
        assert env.frame_skip == {"hammer-v0": 5, "door-v0": 1, "pen-v0": 5, "relocate-v0": 5}[env_id]   # door_v0.py:10: this fork runs door at 1
        m = load_compiled(R.TASK[env_id])
        assert env.frame_skip == m.frame_skip and env.obs_dim == m.obs_dim and env.model.nu == m.nu


@pytest.mark.parametrize("env_id,var", CONFIGS)
def test_restated_task_layer_reproduces_committed_reference_outputs(env_id, var):
    """Runs anywhere: the oracle's own step() (oracle_task.inc) against what the reference's code produced."""
    fx = _load(env_id, var)
    model = load_compiled(R.TASK[env_id])
    if var:
        model = model.with_variation(var)
    E, T = fx["action"].shape[:2]
    ora = OracleSim(model, E, seed=int(fx["seed"]))
    ora.reset()                                                    # zero warm start, counters
    ora.set_state(np.zeros((E, ora.nv)), np.zeros((E, ora.nv)), fx["par"])
    assert np.abs(ora.get_obs() - fx["reset_obs"]).max() <= 1e-7
    for t in range(T):
        obs, rew, done, goal = ora.step(fx["action"][:, t])
        assert np.abs(obs - fx["obs"][:, t]).max() <= 1e-7, t
        assert np.abs(rew - fx["reward"][:, t]).max() <= 1e-6 * max(1.0, np.abs(fx["reward"][:, t]).max()), t   # the ABI's reward is f32
        assert np.array_equal(goal.astype(bool), fx["goal"][:, t]) and np.array_equal(done.astype(bool), fx["done"][:, t]), t
        q, v, _ = ora.get_state()
        assert np.array_equal(q, fx["qpos"][:, t]) and np.array_equal(v, fx["qvel"][:, t])
    from mj_envs_b200.vec_env import evaluate_success
    paths = [dict(env_infos=dict(goal_achieved=fx["goal"][e])) for e in range(E)]
    assert evaluate_success(env_id, paths) == float(fx["success_pct"])
    # thresholds of evaluate_success on synthetic paths: > 25 goal steps (pen: > 20), hammer_v0.py:167-175 / pen_v0.py:180-188
    thr = 20 if env_id == "pen-v0" else 25
    mk = lambda k: dict(env_infos=dict(goal_achieved=np.arange(60) < k))
    assert evaluate_success(env_id, [mk(thr), mk(thr + 1), mk(0), mk(60)]) == 50.0


def test_draw_ranges_recorded_in_the_fixtures():
    for (env_id, var) in CONFIGS:
        fx = _load(env_id, var)
        for ep in range(fx["draws"].shape[0]):
            assert [tuple(x) for x in fx["draws"][ep]] == [tuple(map(float, r)) for r in RANGES[(env_id, var)]]
        assert int(fx["horizon"]) == {"hammer-v0": 200, "door-v0": 200, "pen-v0": 100, "relocate-v0": 200}[env_id]
