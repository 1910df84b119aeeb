"""GPU tests added in build round 2 (VERDICT r1 'Next round' items 5, 7, 8 and ADVICE r1): the multi-stream sub-batch path
against the single-stream one, num_envs = 1, a non-zero S_nail touch reading, fp32 tolerance on every env-step whose
constraint set agrees with the oracle's, the oracle comparison at BASELINE batch sizes, the 33..100-contact class."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TASKS = ["hammer", "door", "pen", "relocate"]


def _rel(x, y):
    return np.abs(x - y).max(axis=-1) / np.maximum(1.0, np.abs(y).max(axis=-1))


def _run_with_subbatches(nsub, task, n, steps, host):
    """A fresh process-wide env var is read by adroit_create, so the two handles of one test see different settings."""
    import torch
    from mj_envs_b200.vec_env import AdroitVecEnv
    old = os.environ.get("ADROIT_SUBBATCHES")
    if nsub is None:
        os.environ.pop("ADROIT_SUBBATCHES", None)
    else:
        os.environ["ADROIT_SUBBATCHES"] = str(nsub)
    try:
        env = AdroitVecEnv(f"{task}-v0", n, seed=5, auto_reset=True)
    finally:
        if old is None:
            os.environ.pop("ADROIT_SUBBATCHES", None)
        else:
            os.environ["ADROIT_SUBBATCHES"] = old
    obs0, _ = env.reset()
    g = torch.Generator(device="cpu"); g.manual_seed(0)
    outs = [obs0.cpu().numpy()]
    obs = np.zeros((n, env.obs_dim), np.float32); rew = np.zeros(n, np.float32); done = np.zeros(n, np.uint8); goal = np.zeros(n, np.uint8)
    for t in range(steps):
        a = torch.rand((n, env.nu), generator=g) * 2 - 1
        if host:
            env.step_host(a.numpy(), obs, rew, done, goal)
            outs += [obs.copy(), rew.copy(), done.copy(), goal.copy()]
        else:
            o, r, d, i = env.step(a)
            outs += [o.cpu().numpy(), r.cpu().numpy(), d.cpu().numpy().astype(np.uint8), i["goal_achieved"].cpu().numpy().astype(np.uint8)]
    q, v, p = env._raw_state()
    outs += [q, v, p, env.get_warmstart()]
    assert env.overflow().sum() == 0
    env.close()
    return outs


@pytest.mark.parametrize("host", [False, True])
def test_sub_batch_pipeline_is_bitwise_equal_to_single_stream(host):
    """every benchmark number comes from the multi-stream sub-batch path (n >= 4096); compare it bit for
    bit with ADROIT_SUBBATCHES=1 through step() and step_host(): observations, rewards, flags and the raw state."""
    n = 8192
    one = _run_with_subbatches(1, "hammer", n, 4, host)
    many = _run_with_subbatches(None, "hammer", n, 4, host)
    assert len(one) == len(many)
    for k, (a, b) in enumerate(zip(one, many)):
        assert np.array_equal(a, b), k


def test_reset_of_8192_envs_equals_two_shards_of_4096():
    import torch
    from mj_envs_b200.vec_env import AdroitVecEnv
    n = 8192
    env = AdroitVecEnv("door-v0", n, seed=9)
    a = AdroitVecEnv("door-v0", n // 2, seed=9, env_id_offset=0)
    b = AdroitVecEnv("door-v0", n // 2, seed=9, env_id_offset=n // 2)
    g, _ = env.reset(); ga, _ = a.reset(); gb, _ = b.reset()
    assert torch.equal(torch.cat([ga, gb]), g)
    act = torch.rand((n, env.nu), device=g.device) * 2 - 1
    for _ in range(2):
        o1 = env.step(act); oa = a.step(act[: n // 2]); ob = b.step(act[n // 2:])
    assert torch.equal(torch.cat([oa[0], ob[0]]), o1[0]) and torch.equal(torch.cat([oa[1], ob[1]]), o1[1])
    for e in (env, a, b):
        e.close()


@pytest.mark.parametrize("task", TASKS)
def test_single_env_reproduces_single_env_behaviour(task):
    """N = 1 must reproduce the reference's single-env behaviour -- the 4-tuple of hammer_v0.py:90 with a
    leading dim of 1, equal to env 0 of a larger batch and to the oracle."""
    import torch
    from mj_envs_b200.vec_env import AdroitVecEnv, make
    from oracle.oracle import OracleSim
    one = make(f"{task}-v0", is_headless=True, variation_type=None, render_mode="rgb_array", width=64, height=64, seed=7, fp64=True)   # helpers.py:59-63
    many = AdroitVecEnv(f"{task}-v0", 8, seed=7, fp64=True)
    ora = OracleSim(task, 1, seed=7)
    assert one.num_envs == 1
    o1, info = one.reset(); om, _ = many.reset(); oo = ora.reset()
    assert info == {} and o1.shape == (1, one.obs_dim) and o1.dtype == torch.float32
    assert torch.equal(o1[0], om[0]) and np.abs(o1.cpu().numpy() - oo).max() < 1e-6
    rng = np.random.default_rng(1)
    for t in range(10):
        a = rng.uniform(-1, 1, (8, one.nu)).astype(np.float32)
        o1, r1, d1, i1 = one.step(torch.from_numpy(a[:1]))
        om, rm, dm, im = many.step(torch.from_numpy(a))
        oo, ro, do_, go = ora.step(a[:1])
        assert o1.shape == (1, one.obs_dim) and r1.shape == (1,) and d1.shape == (1,) and d1.dtype == torch.bool
        assert i1["goal_achieved"].shape == (1,) and i1["goal_achieved"].dtype == torch.bool
        assert torch.equal(o1[0], om[0]) and torch.equal(r1[0], rm[0])
        assert np.abs(o1.cpu().numpy() - oo).max() < 1e-5 and abs(float(r1[0]) - float(ro[0])) < 1e-5 * max(1.0, abs(float(ro[0])))
    one.close(); many.close()


def test_constructor_validates_the_reference_keyword_arguments():
    from mj_envs_b200.vec_env import make
    with pytest.raises(ValueError):
        make("hammer-v0", width=128, height=128)
    with pytest.raises(ValueError):
        make("hammer-v0", render_mode="depth")
    e = make("door-v0", num_envs=2, render_mode="rgb_array", is_headless=True, variation_type="mass")   # door ignores variation_type (door_v0.py:29)
    assert e.render().shape == (2, 64, 64, 3)
    e.close()


def test_touch_sensor_reads_a_hammer_blow_like_the_oracle():
    """Row a2.9: hammer_v0.py:102 feeds sensordata[S_nail] into obs[45].  Random play never hits the nail, so the hammer
    head is placed in front of the nail head and thrown at it: the reading must be non-zero on both sides and agree
    (raw sensor value through get_solver, and the clipped obs[45])."""
    import torch
    from mj_envs_b200.vec_env import AdroitVecEnv
    from mj_envs_b200.mjcf import load_compiled, quat_to_mat
    from oracle.oracle import OracleSim
    n = 128
    m = load_compiled("hammer")
    ora = OracleSim("hammer", n, seed=2)
    ora.reset()
    q, v, p = ora.get_state()
    _, bq, sx = ora.frames()
    J = lambda name: m.name2id("joint", name)
    tool, target, goal, obj = m.name2id("site", "tool"), m.name2id("site", "S_target"), m.name2id("site", "nail_goal"), m.name2id("body", "Object")
    jt = [J("OBJTx"), J("OBJTy"), J("OBJTz")]
    rng = np.random.default_rng(0)
    for e in range(n):
        R = quat_to_mat(bq[e, obj])
        A = np.stack([R @ m.jnt_axis[j] for j in jt], 1)              # world axes of the object's three slide joints
        nout = sx[e, target] - sx[e, goal]; nout /= np.linalg.norm(nout)   # the nail's axis, pointing out of the board
        d = sx[e, target] + nout * rng.uniform(0.0, 0.03) - sx[e, tool] + rng.uniform(-0.01, 0.01, 3)
        q[e, jt] += A.T @ d
        v[e, jt] = A.T @ (-nout * rng.uniform(0.5, 2.0))              # flying into the nail head
    for fp64 in (True, False):
        env = AdroitVecEnv("hammer-v0", n, seed=2, fp64=fp64)
        env.reset()
        o2 = OracleSim("hammer", n, seed=2); o2.reset(); o2.set_state(q, v, p)
        hits, peak = 0, 0.0
        for t in range(6):
            qq, vv, pp = o2.get_state()
            env._set_raw_state(qq, vv, pp); env.set_warmstart(o2.get_warmstart())
            a = np.zeros((n, env.nu), np.float32)
            oo, _, _, _ = o2.step(a)
            go, _, _, _ = env.step(torch.from_numpy(a))
            to, tg = o2.solver()[3], env.solver()[3]                   # raw sensordata[S_nail] of the step's last forward
            hits += int((to > 1e-3).sum()); peak = max(peak, float(to.max()))
            if fp64:
                assert np.abs(tg - to).max() <= 1e-6 * max(1.0, float(to.max())), (t, np.abs(tg - to).max())
                assert np.abs(go.cpu().numpy()[:, 45] - oo[:, 45]).max() <= 1e-6
            else:   # fp32: the blow is a stiff contact; the bulk agrees to 1e-3 of the peak force, stated fraction as in the parity tests
                ok = np.abs(tg - to) <= 2e-3 * max(1.0, float(to.max()))
                assert ok.mean() >= 0.97, (t, ok.mean())
        assert hits > 20 and peak > 1.0, (hits, peak)
        env.close()


def _constraint_sets(env, ora):
    """Per env: does the constraint set of the step's forward agree between the CUDA path and the oracle?  Same number of
    friction / limit rows, same contact list (geom pairs, in order) and the same contact geometry (position 1e-4, distance
    1e-5: an MPR portal flip or a different box feature moves the point by millimetres).  Pure reads on both sides."""
    n = env.num_envs
    L = env.debug_layout()
    gi = np.zeros((n, L["GI_TOTAL"]), dtype=np.int32)
    env.abi.lib.adroit_debug_iscratch(env._h, gi.ctypes.data, None)
    con = env.debug_scratch(L["G_CON"], 16 * L["MAXCON"]).reshape(n, L["MAXCON"], 16)
    _, ncon_o, nefc_o, _ = ora.solver()
    same = np.zeros(n, dtype=bool)
    for e in range(n):
        oc = ora.contacts(e)
        k = int(gi[e, 0])
        if k != len(oc) or k != int(ncon_o[e]):
            continue
        rows = sum(1 if int(d) == 1 else 2 * (int(d) - 1) for d in oc[:, 9])
        if int(gi[e, 1]) != int(nefc_o[e]) - rows:
            continue
        g = gi[e, L["GI_CON"]:L["GI_CON"] + 4 * k].reshape(k, 4)
        if k and (not np.array_equal(g[:, 0], oc[:, 7].astype(int)) or not np.array_equal(g[:, 1], oc[:, 8].astype(int))):
            continue
        if k and (np.abs(con[e, :k, 1:4] - oc[:, 1:4]).max() > 1e-4 or np.abs(con[e, :k, 0] - oc[:, 0]).max() > 1e-5):
            continue
        same[e] = True
    return same


@pytest.mark.parametrize("task", TASKS)
def test_fp32_tolerance_holds_wherever_the_constraint_set_agrees(task):
    """the 1e-4 north-star tolerance is asserted on EVERY env-step whose constraint set (rows, contact
    list, contact geometry) agrees between the fp32 CUDA path and the fp64 oracle, for all four tasks; env-steps where the
    sets differ (a pair crossing `margin`, an MPR portal flip) are the only unbounded ones, and those must be rare.  The scenes are
    run at frame_skip = 1 so that one env-step is one mj_step and the step's own forward is what both sides still hold."""
    import torch
    from mj_envs_b200.vec_env import AdroitVecEnv
    from mj_envs_b200.mjcf import load_compiled, Model
    from oracle.oracle import OracleSim
    n, steps = 96, 60
    m = load_compiled(task)
    m1 = Model(task=m.task, names=m.names, A=dict(m.A)); m1.A["sizes"] = m1.A["sizes"].copy(); m1.A["sizes"][9] = 1
    env = AdroitVecEnv(f"{task}-v0", n, seed=3, model=m1)
    ora = OracleSim(m1, n, seed=3)
    env.reset(); ora.reset()
    rng = np.random.default_rng(3)
    agree = differ = 0
    worst_agree = 0.0
    outliers = []
    within = 0
    for t in range(steps):
        q, v, p = ora.get_state()
        env._set_raw_state(q, v, p); env.set_warmstart(ora.get_warmstart())
        a = rng.uniform(-1, 1, (n, env.nu)).astype(np.float32)
        if t % 8 < 4:
            a[:] = rng.uniform(-1, 1, (1, env.nu)).astype(np.float32)
        ora.step(a)
        env.step(torch.from_numpy(a))
        same = _constraint_sets(env, ora)
        gq, gv, _ = env._raw_state(); nq, nv_, _ = ora.get_state()
        e = np.maximum(_rel(gq, nq), _rel(gv, nv_))
        agree += int(same.sum()); differ += int((~same).sum())
        if same.any():
            worst_agree = max(worst_agree, float(e[same].max()))
        within += int((same & (e < 1e-4)).sum())
        outliers += [(t, int(i), float(e[i])) for i in np.nonzero(same & (e >= 1e-4))[0]]
    assert agree > 0.95 * (agree + differ), (agree, differ)
    # measured on B200 (tools/gpu_fp32_outliers.py, profiles/R2e_fp32_outliers.txt): worst agreeing env-step 1.0e-6 (hammer),
    # 6.6e-6 (door), 4.6e-5 (pen), 1.2e-5 (relocate) -- after the fp32 cost-decrease floor of the Newton stop rule went from 2e-6
    # to 2e-7 (at 2e-6 pen had 4 of 5760 agreeing env-steps at 1.1e-4 .. 4.6e-4).  So: EVERY agreeing env-step inside 1e-4.
    assert not outliers, (outliers[:5], len(outliers), agree, worst_agree)
    assert within == agree
    env.close()


@pytest.mark.parametrize("task,n", [("door", 4096), ("hammer", 8192), ("pen", 16384)])
def test_fp32_single_step_against_the_oracle_at_baseline_batch_sizes(task, n):
    """VERDICT r1 weak #2b: the oracle comparison at BASELINE.json's batch sizes, on every 64th env (the oracle runs the strided
    sample): free-run the CUDA batch into the episode, copy the sampled envs' states into the oracle, step both once."""
    import torch
    from mj_envs_b200.vec_env import AdroitVecEnv
    from oracle.oracle import OracleSim
    env = AdroitVecEnv(f"{task}-v0", n, seed=13)
    env.reset()
    g = torch.Generator(device=env.device); g.manual_seed(1)
    for _ in range(12):
        env.step(torch.rand((n, env.nu), device=env.device, generator=g) * 2 - 1)
    idx = np.arange(0, n, 64)
    q, v, p = env._raw_state(); w = env.get_warmstart()
    ora = OracleSim(task, len(idx), seed=13)
    ora.reset(); ora.set_state(q[idx], v[idx], p[idx]); ora.set_warmstart(w[idx])
    a = (torch.rand((n, env.nu), device=env.device, generator=g) * 2 - 1)
    obs, rew, _, _ = env.step(a)
    oobs, orew, _, _ = ora.step(a.cpu().numpy()[idx])
    gq, gv, _ = env._raw_state(); nq, nv_, _ = ora.get_state()
    eq, ev = _rel(gq[idx], nq), _rel(gv[idx], nv_)
    er = np.abs(rew.cpu().numpy()[idx] - orew) / np.maximum(1.0, np.abs(orew))
    assert np.median(eq) < 1e-6 and np.median(ev) < 1e-5
    assert (eq < 1e-4).mean() >= 0.99 and (ev < 1e-4).mean() >= 0.98 and (er < 1e-4).mean() >= 0.99, ((eq < 1e-4).mean(), (ev < 1e-4).mean(), (er < 1e-4).mean())
    assert env.overflow().sum() == 0
    env.close()
