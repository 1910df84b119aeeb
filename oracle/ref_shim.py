"""Reference-task shim.  TEST INFRASTRUCTURE ONLY.

Runs the reference's OWN task files -- /root/reference/mj_envs_vision/hand_manipulation_suite/{hammer,door,pen,
relocate}_v0.py, imported read-only and UNMODIFIED -- on top of the fp64 CPU oracle: the third-party modules they import
(`mujoco_py`, `mjrl.envs.mujoco_env`, `gym`; none installable here) are replaced in `sys.modules` by the minimal stubs
below, whose `MjSim.step / forward / reset / get_state / set_state` are the oracle's raw `mj_step / mj_forward /
mj_resetData` (oracle_task.inc: oracle_raw_*).  Everything the reference's `step / get_obs / reset_model /
get_env_state / set_env_state / evaluate_success` computes is then the reference's code, not a restatement; comparing it
with oracle_task.inc (and with the CUDA path) pins SURVEY §8 rows a1, a3, a4, a5, a7, a8, a9 to the reference itself.

What is stubbed, and from where its behaviour is restated:
  * mujoco_py.MjSim / load_model_from_path / MjViewer -- the call surface listed in SURVEY.md §8(b) "Downward"
    (hammer_v0.py:29-32,44-50,62-66,96-102,107-125,138-153 and the twins); physics = the oracle (MuJoCo itself is
    unavailable: parity of the PHYSICS stays unpinned, only the task layer is pinned here).
  * mjrl.envs.mujoco_env.MujocoEnv -- [EXT, recalled: mjrl is not under /root/reference] __init__ (one step(zeros) with
    unscaled ctrl), do_simulation (write data.ctrl, frame_skip x sim.step()), set_state (+forward),
    reset (sim.reset, sim.forward, reset_model), seed.
  * gym -- utils.EzPickle, spaces.Box, envs.registration.register (records ids / horizons: row a9), Env.
The writable model arrays the reference pokes (body_pos, body_quat, site_pos, body_mass, geom_pos, geom_size) are mapped
onto the oracle's per-env parameter vector; a write the mapping does not cover raises (so nothing the reference does to
the model can be silently dropped), except writes to the world body's row, which MuJoCo ignores (App. A.8 Q4).
"""
import ctypes as C
import importlib
import os
import sys
import types

import numpy as np

from oracle.oracle import OracleSim, _p, abi
from mj_envs_b200.mjcf import load_compiled

REFERENCE_ROOT = "/root/reference"
REGISTERED = {}          # env id -> dict(entry_point, max_episode_steps), filled by the reference's own register() calls
_XML_TASK = {"DAPG_hammer.xml": "hammer", "DAPG_door.xml": "door", "DAPG_pen.xml": "pen", "DAPG_relocate.xml": "relocate"}


def available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "mj_envs_vision", "hand_manipulation_suite"))


def _raw():
    L = abi().lib
    if not getattr(L, "_raw_ready", False):
        for name in ("oracle_raw_mj_step", "oracle_raw_mj_forward"):
            f = getattr(L, name); f.restype = C.c_int; f.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.oracle_raw_reset_data.restype = C.c_int; L.oracle_raw_reset_data.argtypes = [C.c_void_p, C.c_int]
        L.oracle_raw_poke.restype = C.c_int; L.oracle_raw_poke.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 3
        L.oracle_debug_uniforms.restype = None
        L.oracle_debug_uniforms.argtypes = [C.c_uint64, C.c_uint64, C.c_uint32, C.c_void_p]
        L._raw_ready = True
    return L


def philox_uniforms(seed, global_env, episode):
    """The uniforms the oracle / CUDA reset draws for (seed, global env, episode), in draw order."""
    out = np.zeros(8)
    _raw().oracle_debug_uniforms(int(seed), int(global_env), int(episode), _p(out))
    return out


# ----------------------------------------------------------------------------------------------- mujoco_py stub
class _Opt:
    def __init__(self, timestep):
        self.timestep = timestep


class ShimModel:
    """mujoco_py.PyMjModel surface used by the reference."""

    def __init__(self, compiled):
        m = self._m = compiled
        self.nq = self.nv = m.nv
        self.nu = m.nu
        self.opt = _Opt(float(m.opt_real[0]))
        self.body_pos = m.body_pos.copy()
        self.body_quat = m.body_quat.copy()
        self.site_pos = m.site_pos.copy()
        self.body_mass = m.body_mass.copy()
        self.geom_pos = m.geom_pos.copy()
        self.geom_size = m.geom_size.copy()
        self.geom_rgba = m.geom_rgba.copy()
        self.actuator_ctrlrange = m.act_ctrlrange.copy()
        self.jnt_dofadr = np.arange(m.nv, dtype=np.int32)          # every joint is a 1-dof hinge / slide
        # general actuators, affine bias (DAPG_assets.xml:243-266): gainprm[0], biasprm[0:3]; 10 columns as in MuJoCo 2.1
        self.actuator_gainprm = np.zeros((m.nu, 10)); self.actuator_gainprm[:, 0] = m.act_gain
        self.actuator_biasprm = np.zeros((m.nu, 10)); self.actuator_biasprm[:, :3] = m.act_bias
        self._defaults = {k: getattr(self, k).copy() for k in
                          ("body_pos", "body_quat", "site_pos", "body_mass", "geom_pos", "geom_size", "actuator_gainprm", "actuator_biasprm")}

    def _id(self, kind, name):
        try:
            return self._m.names[kind].index(name)
        except ValueError:
            raise ValueError(f'No "{kind}" with name {name} exists.')

    def body_name2id(self, n): return self._id("body", n)
    def site_name2id(self, n): return self._id("site", n)
    def geom_name2id(self, n): return self._id("geom", n)
    def joint_name2id(self, n): return self._id("joint", n)
    def actuator_name2id(self, n): return self._id("actuator", n)
    def sensor_name2id(self, n): return self._id("sensor", n)


class ShimData:
    def __init__(self, m):
        self.qpos = np.zeros(m.nv); self.qvel = np.zeros(m.nv); self.ctrl = np.zeros(m.nu)
        self.body_xpos = np.zeros((m.nbody, 3)); self.body_xquat = np.zeros((m.nbody, 4)); self.site_xpos = np.zeros((m.nsite, 3))
        self.sensordata = np.zeros(len(m.names["sensor"]))   # every sensor is 1-D, S_nail is sensor 0 (App. A.8 Q10)
        self.cam_xpos = m.cam_pos.copy() if "cam_pos" in m.A else np.zeros((1, 3))
        self.time = 0.0


class MjSimState:
    def __init__(self, time, qpos, qvel, act=None, udd_state=None):
        self.time, self.qpos, self.qvel, self.act, self.udd_state = time, qpos, qvel, act, udd_state or {}


class MjSim:
    """mujoco_py.MjSim backed by one env of the fp64 oracle."""

    def __init__(self, model):
        self.model = model
        self._compiled = model._m
        self._variation = None
        self._compiled_var = self._compiled
        self._ora = OracleSim(self._compiled, 1)
        self.data = ShimData(self._compiled)
        self.nsteps = 0
        self.forward()

    # --- model writes -> the oracle's per-env parameters
    def _sync_model(self):
        M, cm, d0 = self.model, self._compiled, self.model._defaults
        for k in ("actuator_gainprm", "actuator_biasprm"):   # App. A.8 Q3: the reference's overwrite equals the XML values
            if not np.array_equal(getattr(M, k), d0[k]):
                raise RuntimeError(f"reference changed model.{k}; the compiled scene would no longer match")
        task = cm.task
        var = None
        if task == "hammer":
            obj, head, neck = cm.name2id("body", "Object"), cm.name2id("geom", "head"), cm.name2id("geom", "neck")
            dm = M.body_mass != d0["body_mass"]; dp = M.geom_pos != d0["geom_pos"]; dz = M.geom_size != d0["geom_size"]
            if dm.any(): var = "mass"
            if dp.any(): var = "pos" if var is None else "mixed"
            if dz.any(): var = "size" if var is None else "mixed"
            if var == "mixed":
                raise RuntimeError("more than one hammer variation written at once")
            ok_m = np.zeros_like(dm); ok_m[obj] = True
            ok_p = np.zeros_like(dp); ok_p[head, 0] = ok_p[neck, 0] = True
            ok_z = np.zeros_like(dz); ok_z[head, 0] = ok_z[head, 1] = True
            if (dm & ~ok_m).any() or (dp & ~ok_p).any() or (dz & ~ok_z).any():
                raise RuntimeError("reference wrote a model field the oracle has no parameter for")
        else:
            for k in ("body_mass", "geom_pos", "geom_size"):
                if not np.array_equal(getattr(M, k), d0[k]):
                    raise RuntimeError(f"reference wrote model.{k} in task {task}")
        if var != self._variation:          # first write of a variation field: switch to the scene compiled with that variation
            q, v, _ = self._ora.get_state(); w = self._ora.get_warmstart()
            self._variation = var
            self._compiled_var = cm.with_variation(var) if var else cm
            self._ora = OracleSim(self._compiled_var, 1)
            _raw().oracle_raw_poke(self._ora.h, 0, _p(q), _p(v), None)
            self._ora.set_warmstart(w)
        cv = self._compiled_var if self._variation else cm
        par = np.array(cv.par0, dtype=np.float64).copy()
        bp, sp = cm.body_par, cm.site_par
        for b in range(cm.nbody):
            for col, arr, w in ((0, M.body_pos, 3), (1, M.body_quat, 4)):
                off = bp[b, col]
                if off >= 0:
                    par[off:off + w] = arr[b]
                elif b > 0 and not np.array_equal(arr[b], (d0["body_pos"], d0["body_quat"])[col][b]):   # row 0 = world: ignored by MuJoCo
                    raise RuntimeError(f"reference wrote body {b} {('pos', 'quat')[col]} which is not a per-env parameter")
        for s in range(cm.nsite):
            if sp[s] >= 0:
                par[sp[s]:sp[s] + 3] = M.site_pos[s]
            elif not np.array_equal(M.site_pos[s], d0["site_pos"][s]):
                raise RuntimeError(f"reference wrote site {s} pos which is not a per-env parameter")
        if var == "mass": par[3] = M.body_mass[obj]
        elif var == "pos": par[3], par[4] = M.geom_pos[head, 0], M.geom_pos[neck, 0]
        elif var == "size": par[3], par[4] = M.geom_size[head, 0], M.geom_size[head, 1]
        return par

    def _push(self):
        par = self._sync_model()
        q = np.ascontiguousarray(self.data.qpos, dtype=np.float64); v = np.ascontiguousarray(self.data.qvel, dtype=np.float64)
        _raw().oracle_raw_poke(self._ora.h, 0, _p(q), _p(v), _p(par))

    def _pull(self):
        q, v, _ = self._ora.get_state()
        self.data.qpos[:] = q[0]; self.data.qvel[:] = v[0]
        bx, bq, sx = self._ora.frames()
        self.data.body_xpos[:] = bx[0]; self.data.body_xquat[:] = bq[0]; self.data.site_xpos[:] = sx[0]
        self.data.sensordata[:] = 0.0
        self.data.sensordata[0] = self._ora.solver()[3][0]      # S_nail touch (DAPG_hammer.xml:34-36); 0 elsewhere

    # --- mujoco_py.MjSim API
    def step(self):
        self._push()
        c = np.ascontiguousarray(self.data.ctrl, dtype=np.float64)
        _raw().oracle_raw_mj_step(self._ora.h, 0, _p(c))
        self.data.time += self.model.opt.timestep
        self.nsteps += 1
        self._pull()

    def forward(self):
        self._push()
        c = np.ascontiguousarray(self.data.ctrl, dtype=np.float64)
        _raw().oracle_raw_mj_forward(self._ora.h, 0, _p(c))
        self._pull()

    def reset(self):
        _raw().oracle_raw_reset_data(self._ora.h, 0)
        self.data.qpos[:] = 0; self.data.qvel[:] = 0; self.data.ctrl[:] = 0; self.data.time = 0.0

    def get_state(self):
        return MjSimState(self.data.time, self.data.qpos.copy(), self.data.qvel.copy(), np.zeros(0))   # na = 0: act empty

    def set_state(self, st):
        self.data.time = st.time; self.data.qpos[:] = st.qpos; self.data.qvel[:] = st.qvel

    def render(self, *a, **k):
        raise RuntimeError("GL rendering: unavailable on this box")

    # test helpers (not part of mujoco_py)
    def _warmstart(self): return self._ora.get_warmstart()[0]
    def _set_warmstart(self, w): self._ora.set_warmstart(np.asarray(w, dtype=np.float64).reshape(1, -1))
    def _par(self): return self._ora.get_state()[2][0]


def load_model_from_path(path):
    return ShimModel(load_compiled(_XML_TASK[os.path.basename(path)]))


class _Cam:
    azimuth = 0.0; distance = 0.0; elevation = 0.0


class MjViewer:
    def __init__(self, sim):
        self.sim, self.cam = sim, _Cam()

    def render(self):
        pass


# ----------------------------------------------------------------------------------------------- gym / mjrl stubs
class EzPickle:
    def __init__(self, *a, **k):
        pass


class Box:
    def __init__(self, low, high, dtype=np.float32, shape=None):
        self.low, self.high, self.dtype = np.asarray(low, dtype=dtype), np.asarray(high, dtype=dtype), dtype
        self.shape = self.low.shape


class Env:
    metadata = {}


class RecordingRandom:
    """np_random stand-in: `uniform(low, high)` returns low + u*(high-low) for the next u of a supplied stream (so the
    reference's draw ORDER and RANGES can be checked against the Philox stream of the CUDA / oracle reset) and logs calls."""

    def __init__(self, seed=None):
        self._gen = np.random.default_rng(seed)
        self.queue, self.calls = [], []

    def feed(self, us):
        self.queue = list(us)

    def uniform(self, low=0.0, high=1.0, size=None):
        assert size is None
        u = self.queue.pop(0) if self.queue else self._gen.random()
        self.calls.append((low, high, u))
        return low + u * (high - low)


def _register(id, entry_point=None, max_episode_steps=None, **kw):
    REGISTERED[id] = dict(entry_point=entry_point, max_episode_steps=max_episode_steps, **kw)


class MujocoEnv(Env):
    """[EXT] mjrl.envs.mujoco_env.MujocoEnv, restated (mjrl is a git dependency, Dockerfile:80, absent from /root/reference)."""

    def __init__(self, model_path=None, frame_skip=1, sim=None):
        self.sim = sim if sim is not None else MjSim(load_model_from_path(model_path))
        self.data = self.sim.data
        self.model = self.sim.model
        self.frame_skip = frame_skip
        self.metadata = {"render.modes": ["human", "rgb_array"], "video.frames_per_second": int(np.round(1.0 / self.dt))}
        self.mujoco_render_frames = False
        self.init_qpos = self.data.qpos.ravel().copy()
        self.init_qvel = self.data.qvel.ravel().copy()
        observation, _reward, done, _info = self.step(np.zeros(self.model.nu))     # App. A.8 Q5
        assert not done
        self.obs_dim = int(np.sum([o.size for o in observation])) if type(observation) is tuple else observation.size
        bounds = self.model.actuator_ctrlrange.copy()
        self.action_space = Box(bounds[:, 0], bounds[:, 1], dtype=np.float32)
        high = np.inf * np.ones(self.obs_dim)
        self.observation_space = Box(-high, high, dtype=np.float32)
        self.seed()

    def seed(self, seed=None):
        self.np_random = RecordingRandom(seed)
        return [seed]

    @property
    def dt(self):
        return self.model.opt.timestep * self.frame_skip

    def reset(self, seed=None):
        self.sim.reset()
        self.sim.forward()
        return self.reset_model()

    def set_state(self, qpos, qvel):
        assert qpos.shape == (self.model.nq,) and qvel.shape == (self.model.nv,)
        state = self.sim.get_state()
        for i in range(self.model.nq): state.qpos[i] = qpos[i]
        for i in range(self.model.nv): state.qvel[i] = qvel[i]
        self.sim.set_state(state)
        self.sim.forward()

    def do_simulation(self, ctrl, n_frames):
        for i in range(self.model.nu):
            self.sim.data.ctrl[i] = ctrl[i]
        for _ in range(n_frames):
            self.sim.step()
            if self.mujoco_render_frames is True:
                self.mj_render()

    def mj_render(self):
        pass


def _module(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    return m


_INSTALLED = None


def install_stubs():
    """Put the stub modules into sys.modules (only those that are really missing for cv2 / torchvision) and /root/reference on sys.path."""
    global _INSTALLED
    if _INSTALLED:
        return
    g_utils = _module("gym.utils", EzPickle=EzPickle)
    g_spaces = _module("gym.spaces", Box=Box)
    g_reg = _module("gym.envs.registration", register=_register)
    g_envs = _module("gym.envs", registration=g_reg)
    gym = _module("gym", utils=g_utils, spaces=g_spaces, envs=g_envs, Env=Env)
    mj = _module("mujoco_py", MjSim=MjSim, MjViewer=MjViewer, load_model_from_path=load_model_from_path, MjSimState=MjSimState,
                 MjRenderContextOffscreen=object)
    me = _module("mjrl.envs.mujoco_env", MujocoEnv=MujocoEnv)
    mjrl_envs = _module("mjrl.envs", mujoco_env=me)
    mjrl = _module("mjrl", envs=mjrl_envs)
    mods = {"gym": gym, "gym.utils": g_utils, "gym.spaces": g_spaces, "gym.envs": g_envs, "gym.envs.registration": g_reg,
            "mujoco_py": mj, "mjrl": mjrl, "mjrl.envs": mjrl_envs, "mjrl.envs.mujoco_env": me}
    for extra in ("cv2", "torchvision"):             # imported (not used on this path) by headless_observer.py
        try:
            importlib.import_module(extra)
        except Exception:
            t = _module(extra)
            if extra == "torchvision":
                tr = _module("torchvision.transforms", Resize=lambda *a, **k: None, CenterCrop=lambda *a, **k: None)
                t.transforms = tr
                mods["torchvision.transforms"] = tr
            mods[extra] = t
    for k, v in mods.items():
        if k in sys.modules and k.split(".")[0] in ("gym", "mujoco_py", "mjrl"):
            raise RuntimeError(f"a real {k} is already imported; the shim is for boxes without it")
        sys.modules[k] = v
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    _INSTALLED = True


ENV_CLASSES = {"hammer-v0": "HammerEnvV0", "door-v0": "DoorEnvV0", "pen-v0": "PenEnvV0", "relocate-v0": "RelocateEnvV0"}


def make_reference_env(env_id, variation_type=None):
    """gym.make(env_id) equivalent: the reference's own class, constructed as helpers.py:59-63 does (no GL: is_headless=False
    takes the MjViewer branch, whose stub holds the camera attributes and draws nothing)."""
    install_stubs()
    suite = importlib.import_module("mj_envs_vision.hand_manipulation_suite")      # runs the reference's register() calls too
    cls = getattr(suite, ENV_CLASSES[env_id])
    return cls(render_mode="rgb_array", width=64, height=64, is_headless=False, variation_type=variation_type)


# ----------------------------------------------------------------------------------------------- the comparison
N_DRAWS = {("hammer-v0", None): 1, ("hammer-v0", "mass"): 2, ("hammer-v0", "pos"): 2, ("hammer-v0", "size"): 3,
           ("door-v0", None): 3, ("pen-v0", None): 2, ("relocate-v0", None): 5}
TASK = {"hammer-v0": "hammer", "door-v0": "door", "pen-v0": "pen", "relocate-v0": "relocate"}


def _restated_reward(ora):
    L = _raw()
    if not hasattr(L, "_rew_ready"):
        L.oracle_debug_reward.restype = C.c_int
        L.oracle_debug_reward.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L._rew_ready = True
    r = np.zeros(1); g = np.zeros(1, np.int32); d = np.zeros(1, np.int32)
    L.oracle_debug_reward(ora.h, 0, _p(r), _p(g), _p(d))
    return float(r[0]), int(g[0]), int(d[0])


def state_dict_to_par(task, sd, par):
    """The per-env parameters a reference state dict implies (what set_env_state writes into the model)."""
    par = np.array(par, dtype=np.float64).copy()
    if task == "hammer": par[:3] = sd["board_pos"]            # hammer_v0.py:144-153
    elif task == "door": par[:3] = sd["door_body_pos"]        # door_v0.py:130-138
    elif task == "pen": par[:4] = sd["desired_orien"]         # pen_v0.py:143-152
    else: par[:3] = sd["obj_pos"]; par[3:6] = sd["target_pos"]   # relocate_v0.py:118-129
    return par


def rollout_against_restatement(env_id, variation_type=None, seed=3, episodes=4, steps=50, action_scale=1.25):
    """Run the reference's own env class (on the shim) next to the oracle's restated task layer (oracle_task.inc) from the
    same seeds and actions.  Returns (fixture dict of arrays, report dict of max deviations).  Physics is the oracle on
    both sides, so every deviation is a task-layer deviation."""
    task = TASK[env_id]
    ref = make_reference_env(env_id, variation_type)
    ref2 = make_reference_env(env_id, variation_type)      # receives set_env_state
    model = load_compiled(task)
    if variation_type:
        model = model.with_variation(variation_type)
    ora = OracleSim(model, 1, seed=seed)
    ora2 = OracleSim(model, 1, seed=seed)
    rng = np.random.default_rng(seed + 1000)
    nu, nv = ora.nu, ora.nv
    F = dict(action=[], obs=[], reward=[], goal=[], done=[], reset_obs=[], par=[], draws=[], sd_obs=[], qpos=[], qvel=[])
    sd_keys = None
    sds = {}
    rep = dict(par=0.0, reset_obs=0.0, obs=0.0, reward=0.0, state=0.0, flags=0, sd=0.0, sd_obs=0.0, obs_exact=0, obs_total=0)
    for ep in range(episodes):
        us = philox_uniforms(seed, 0, ep)
        ref.np_random.feed(us)
        n0 = len(ref.np_random.calls)
        robs, info = ref.reset()
        assert info == {}
        draws = ref.np_random.calls[n0:]
        assert len(draws) == N_DRAWS[(env_id, variation_type)], (env_id, variation_type, draws)
        oobs = ora.reset()[0]
        rpar, opar = ref.sim._par(), ora.get_state()[2][0]
        rep["par"] = max(rep["par"], float(np.abs(rpar - opar).max()))
        rep["reset_obs"] = max(rep["reset_obs"], float(np.abs(robs.astype(np.float64) - oobs).max()))
        # from here on both sides carry bit-identical parameters (numpy's low + u*(high-low) and the C draw differ in the last bit)
        ora.set_state(np.zeros((1, nv)), np.zeros((1, nv)), rpar[None])
        F["par"].append(rpar.copy()); F["draws"].append([(lo, hi) for lo, hi, _ in draws]); F["reset_obs"].append(robs.copy())
        ea, eo, er, eg, ed, eq, ev = [], [], [], [], [], [], []
        for t in range(steps):
            a = rng.uniform(-action_scale, action_scale, nu).astype(np.float32)     # beyond [-1,1]: exercises the clip
            ro, rr, rd, ri = ref.step(a)
            oo, _, od, og = ora.step(a)
            orr, og64, od64 = _restated_reward(ora)
            q, v, _ = ora.get_state()
            rep["state"] = max(rep["state"], float(np.abs(ref.sim.data.qpos - q[0]).max()), float(np.abs(ref.sim.data.qvel - v[0]).max()))
            rep["obs"] = max(rep["obs"], float(np.abs(ro.astype(np.float64) - oo[0]).max()))
            rep["obs_exact"] += int((ro == oo[0]).sum()); rep["obs_total"] += ro.size
            rep["reward"] = max(rep["reward"], abs(rr - orr) / max(1.0, abs(orr)))
            rep["flags"] += int(bool(ri["goal_achieved"]) != bool(og[0])) + int(bool(rd) != bool(od[0]))
            assert ro.dtype == np.float32 and isinstance(ri["goal_achieved"], bool)
            ea.append(a); eo.append(ro); er.append(rr); eg.append(ri["goal_achieved"]); ed.append(rd); eq.append(q[0]); ev.append(v[0])
        for k, x in (("action", ea), ("obs", eo), ("reward", er), ("goal", eg), ("done", ed), ("qpos", eq), ("qvel", ev)):
            F[k].append(np.array(x))
        # get_env_state / set_env_state: the reference's dict, and the observation a second env shows after loading it
        sd = ref.get_env_state()
        sd_keys = sorted(sd)
        for k in sd:
            sds.setdefault(k, []).append(np.array(sd[k], dtype=np.float64))
        q, v, p = ora.get_state()
        bx, _, sx = ora.frames()
        mine = dict(qpos=q[0], qvel=v[0])
        if task == "hammer": mine.update(board_pos=p[0, :3], target_pos=sx[0, model.name2id("site", "S_target")])
        elif task == "door": mine.update(door_body_pos=p[0, :3])
        elif task == "pen": mine.update(desired_orien=p[0, :4])
        else: mine.update(hand_qpos=q[0, :30], obj_pos=bx[0, model.name2id("body", "Object")], target_pos=sx[0, model.name2id("site", "target")],
                          palm_pos=sx[0, model.name2id("site", "S_grasp")])
        assert sorted(mine) == sd_keys, (sorted(mine), sd_keys)
        rep["sd"] = max([rep["sd"]] + [float(np.abs(np.asarray(sd[k]) - mine[k]).max()) for k in sd])
        ref2.set_env_state(sd)
        o2 = ref2.get_obs()
        ora2.set_state(np.asarray(sd["qpos"])[None], np.asarray(sd["qvel"])[None], state_dict_to_par(task, sd, ora2.get_state()[2][0])[None])
        rep["sd_obs"] = max(rep["sd_obs"], float(np.abs(o2.astype(np.float64) - ora2.get_obs()[0]).max()))
        F["sd_obs"].append(o2.copy())
    # evaluate_success on the recorded paths: the reference's method
    paths = [dict(env_infos=dict(goal_achieved=g)) for g in F["goal"]]
    fix = {k: np.array(v) for k, v in F.items()}
    fix["success_pct"] = np.array(ref.evaluate_success(paths))
    fix["seed"] = np.array(seed)
    for k, v in sds.items():
        fix["sd_" + k] = np.array(v)
    fix["horizon"] = np.array(REGISTERED[env_id]["max_episode_steps"])
    fix["action_low"] = ref.action_space.low.copy(); fix["action_high"] = ref.action_space.high.copy()
    return fix, rep


def crafted_state_dicts(env_id, ref):
    """Reference-format state dicts that put each task into its bonus / goal / done branches (random play rarely gets
    there): door past 0.2 / 1.0 / 1.35 rad; hammer with the nail driven in and with the hammer lifted; pen aligned with the
    target and pen below the drop height; relocate with the object lifted next to the target."""
    task = TASK[env_id]
    ref.np_random.feed(philox_uniforms(0, 0, 0))       # reproducible reset draws
    ref.reset()
    sd0 = {k: np.array(v, dtype=np.float64) for k, v in ref.get_env_state().items()}
    J = lambda n: ref.model.joint_name2id(n)
    out = []

    def variant(**edits):
        sd = {k: v.copy() for k, v in sd0.items()}
        for k, v in edits.items():
            if k.startswith("q_"):
                sd["qpos"][J(k[2:])] = v
            else:
                sd[k] = np.array(v, dtype=np.float64)
        if "hand_qpos" in sd:
            sd["hand_qpos"] = sd["qpos"][:30].copy()
        out.append(sd)

    if task == "door":
        for ang in (0.25, 1.1, 1.4):
            variant(q_door_hinge=ang)
    elif task == "hammer":
        variant(q_nail_dir=0.085)                      # target within 0.01 of the goal: +25, +75, goal_achieved
        variant(q_nail_dir=0.075)                      # within 0.02 only: +25
        variant(q_OBJTz=0.15)                          # hammer (and its head) above 0.04: +2
    elif task == "pen":
        # the pen turned onto the (reset-drawn) desired orientation: search the object's three rotation joints at the reset pose
        r = np.random.default_rng(0)
        best, best_e = -2.0, None
        for _ in range(600):
            e = r.uniform(-1.5, 1.5, 3) if best < 0.9 else best_e + r.normal(0, 0.05, 3)
            sd = {k: v.copy() for k, v in sd0.items()}
            for n, x in zip(("OBJRx", "OBJRy", "OBJRz"), e):
                sd["qpos"][J(n)] = x
            ref.set_env_state(sd)
            o = ref.get_obs().astype(np.float64)
            s_ = float(o[33:36] @ o[36:39])
            if s_ > best:
                best, best_e = s_, e.copy()
        variant(q_OBJRx=best_e[0], q_OBJRy=best_e[1], q_OBJRz=best_e[2])   # similarity > 0.95 at distance ~0: +10, +50, goal_achieved
        variant(q_OBJTz=-0.19, q_OBJTx=0.25)           # pen below 0.075: -5, done
    else:
        obj = sd0["obj_pos"]
        variant(q_OBJTz=0.12, target_pos=[obj[0], obj[1], obj[2] + 0.12 + 0.02])     # lifted, within 0.05 of the target
        variant(q_OBJTz=0.12, target_pos=[obj[0] + 0.08, obj[1], obj[2] + 0.12])     # lifted, within 0.1 only
        variant(q_OBJTz=0.12)                                                        # lifted, far from the target
    return out


def crafted_against_restatement(env_id):
    """Every crafted state: set_env_state on the reference env and on the oracle, one zero-action step on both, compare."""
    task = TASK[env_id]
    ref = make_reference_env(env_id)
    model = load_compiled(task)
    ora = OracleSim(model, 1)
    sds = crafted_state_dicts(env_id, ref)
    F = dict(obs0=[], obs=[], reward=[], goal=[], done=[])
    rep = dict(obs=0.0, reward=0.0, flags=0, goals=0, dones=0, rmax=-1e9)
    for sd in sds:
        ref.sim.reset()                                 # mj_resetData: zero warm start on both sides
        ref.set_env_state(sd)
        o0 = ref.get_obs()
        _raw().oracle_raw_reset_data(ora.h, 0)
        ora.set_state(sd["qpos"][None], sd["qvel"][None], state_dict_to_par(task, sd, model.par0)[None])
        rep["obs"] = max(rep["obs"], float(np.abs(o0.astype(np.float64) - ora.get_obs()[0]).max()))
        a = np.zeros(ora.nu, dtype=np.float32)
        ro, rr, rd, ri = ref.step(a)
        oo, _, od, og = ora.step(a)
        orr, _, _ = _restated_reward(ora)
        rep["obs"] = max(rep["obs"], float(np.abs(ro.astype(np.float64) - oo[0]).max()))
        rep["reward"] = max(rep["reward"], abs(rr - orr) / max(1.0, abs(orr)))
        rep["flags"] += int(bool(ri["goal_achieved"]) != bool(og[0])) + int(bool(rd) != bool(od[0]))
        rep["goals"] += int(ri["goal_achieved"]); rep["dones"] += int(rd); rep["rmax"] = max(rep["rmax"], rr)
        F["obs0"].append(o0); F["obs"].append(ro); F["reward"].append(rr); F["goal"].append(ri["goal_achieved"]); F["done"].append(rd)
    fix = {"crafted_" + k: np.array(v) for k, v in F.items()}
    for k in sds[0]:
        fix["crafted_sd_" + k] = np.array([sd[k] for sd in sds])
    return fix, rep
