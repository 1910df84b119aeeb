"""Hammer variation types (reference hammer_v0.py:110-129): reset_model draws a new hammer mass, head /
neck position or head size into the MODEL of every env without recompiling it (inertia tensor, invweight0, rbound
stay stale).  CPU tests pin the draws of the oracle to the reference's ranges; GPU tests compare the CUDA path with it."""
import numpy as np
import pytest

from mj_envs_b200.mjcf import load_compiled

VARS = ["mass", "pos", "size"]


def test_variation_draws_follow_the_reference_ranges(oracle_mod):
    m = load_compiled("hammer")
    for v in VARS:
        o = oracle_mod.OracleSim(m.with_variation(v), 256, seed=5)
        o.reset()
        p = o.get_state()[2]
        assert p.shape[1] == 5
        assert (p[:, 2] >= 0.1).all() and (p[:, 2] <= 0.25).all()            # board height, hammer_v0.py:109
        if v == "mass":
            assert (p[:, 3] >= 0.05).all() and (p[:, 3] <= 2.5).all() and p[:, 3].std() > 0.3   # :116
        elif v == "pos":
            assert (p[:, 3] >= -0.24).all() and (p[:, 3] <= -0.10).all()      # :120
            assert np.allclose(p[:, 4], -0.14 - (-0.24 - p[:, 3]))            # :122
        else:
            assert (p[:, 3] >= 0.01).all() and (p[:, 3] <= 0.04).all()        # :124
            assert (p[:, 4] >= 0.02).all() and (p[:, 4] <= 0.08).all()        # :125
        # a second reset draws new values (the reference redraws on every reset_model)
        o.reset()
        assert np.abs(o.get_state()[2][:, 3] - p[:, 3]).max() > 1e-3


def test_unknown_variation_raises_and_other_tasks_ignore_it():
    with pytest.raises(Exception, match="Unsupported variation type"):     # hammer_v0.py:129
        load_compiled("hammer").with_variation("colour")
    with pytest.raises(Exception):
        load_compiled("door").with_variation("mass")


def test_mass_variation_changes_the_dynamics_not_the_kinematics(oracle_mod):
    """A heavier hammer falls the same way (free fall is mass independent) but presses harder on the table."""
    m = load_compiled("hammer").with_variation("mass")
    o = oracle_mod.OracleSim(m, 2, seed=0)
    o.reset()
    q, v, p = o.get_state()
    p[0, 3], p[1, 3] = 0.1, 2.0
    o.set_state(q, v, p)
    for _ in range(10):
        o.step(np.zeros((2, o.nu), dtype=np.float32))
    bx, _, _ = o.frames()
    obj = m.name2id("body", "Object")
    assert abs(bx[0, obj, 2] - bx[1, obj, 2]) < 2e-3     # both rest on the table
    _, ncon, _, _ = o.solver()
    assert (ncon > 0).all()


@pytest.mark.gpu
@pytest.mark.parametrize("variation", VARS)
def test_variations_cuda_matches_oracle(variation):
    import torch
    from mj_envs_b200.vec_env import AdroitVecEnv
    from oracle.oracle import OracleSim
    n = 32
    env = AdroitVecEnv("hammer-v0", n, seed=6, fp64=True, variation_type=variation)
    ora = OracleSim(env.model, n, seed=6)
    g, _ = env.reset()
    o = ora.reset()
    assert np.abs(g.cpu().numpy() - o).max() < 2e-6
    assert np.abs(env._raw_state()[2] - ora.get_state()[2]).max() < 1e-12
    rng = np.random.default_rng(1)
    for t in range(20):
        q, v, p = ora.get_state()
        env._set_raw_state(q, v, p)
        env.set_warmstart(ora.get_warmstart())
        a = rng.uniform(-1, 1, (1 if t % 4 < 2 else n, env.nu)).astype(np.float32) * np.ones((n, 1), dtype=np.float32)
        oo, orew, _, _ = ora.step(a)
        go, grew, _, _ = env.step(torch.from_numpy(a))
        gq, gv, _ = env._raw_state()
        nq, nv_, _ = ora.get_state()
        assert np.abs(gq - nq).max() < 1e-9 and np.abs(gv - nv_).max() < 1e-7, (variation, t)
        assert np.abs(grew.cpu().numpy() - orew).max() < 1e-4
    assert env.overflow().sum() == 0
    env.close()


@pytest.mark.gpu
def test_variation_is_ignored_by_the_other_tasks():
    import torch
    from mj_envs_b200.vec_env import AdroitVecEnv
    a = AdroitVecEnv("door-v0", 8, seed=2, variation_type="mass")     # door_v0.py:29 stores it and never reads it
    b = AdroitVecEnv("door-v0", 8, seed=2)
    assert torch.equal(a.reset()[0], b.reset()[0])
    a.close(); b.close()
