"""AdroitVecEnv — batched drop-in for the reference's four gym environments.

Mirrors the per-env API of mj_envs_vision/hand_manipulation_suite/{door,hammer,pen,relocate}_v0.py with every
quantity batched on dim 0: `reset() -> (obs[N,obs_dim] f32, {})`, `step(a[N,nu]) -> (obs, reward[N],
done[N], {"goal_achieved": bool[N]})` (old gym 4-tuple, hammer_v0.py:90), `get_obs()`, `get_env_state()` /
`set_env_state()` with the reference's dict keys, `evaluate_success(paths)` with the reference's
thresholds, env ids and horizons of mj_envs_vision/__init__.py:4-29.

All compute runs in libadroit_b200.so (hand-written sm_100a CUDA behind the C ABI of include/adroit_b200.h); torch is
used for device buffers and streams only.  There is no CPU fallback: construction fails loudly without the library or
without a CUDA device.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import torch

from . import _cabi
from .mjcf import Model, load_compiled

ENV_IDS = {"door-v0": "door", "hammer-v0": "hammer", "pen-v0": "pen", "relocate-v0": "relocate"}
MAX_EPISODE_STEPS = {"door-v0": 200, "hammer-v0": 200, "pen-v0": 100, "relocate-v0": 200}  # mj_envs_vision/__init__.py:4-29

_LIB_PATH = os.environ.get("ADROIT_B200_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "libadroit_b200.so")
_abi = None


def cuda_abi() -> _cabi.CABI:
    global _abi
    if _abi is None:
        if not os.path.exists(_LIB_PATH):
            raise RuntimeError(f"{_LIB_PATH} is missing: build it with `python -m mj_envs_b200.build` "
                               "(the CUDA path has no CPU fallback)")
        _abi = _cabi.CABI(_LIB_PATH, "adroit")
        L = _abi.lib
        L.adroit_debug_scratch.restype = C.c_int
        L.adroit_debug_scratch.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.adroit_debug_iscratch.restype = C.c_int
        L.adroit_debug_iscratch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.adroit_debug_layout.restype = C.c_int
        L.adroit_debug_layout.argtypes = [C.c_void_p, C.c_int]
    return _abi


class Box:
    """Minimal stand-in for gym.spaces.Box (gym is not a dependency)."""

    def __init__(self, low, high, shape, dtype=np.float32):
        self.low = np.full(shape, low, dtype=dtype)
        self.high = np.full(shape, high, dtype=dtype)
        self.shape = tuple(shape)
        self.dtype = dtype

    def sample(self, rng=None):
        rng = rng or np.random.default_rng()
        return rng.uniform(-1, 1, self.shape).astype(self.dtype)


def evaluate_success(env_id: str, paths) -> float:
    """hammer_v0.py:167-175 / door_v0.py:149-157 / relocate_v0.py:141-149 (> 25 goal steps); pen_v0.py:180-188 (> 20)."""
    thr = 20 if env_id == "pen-v0" else 25
    num_success = sum(1 for path in paths if np.sum(np.asarray(path["env_infos"]["goal_achieved"])) > thr)
    return num_success * 100.0 / len(paths)


class AdroitVecEnv:
    """`render_mode, width, height, is_headless, variation_type` are the reference constructors' arguments
    (hammer_v0.py:12, door_v0.py:15, pen_v0.py:12, relocate_v0.py:11; passed by helpers.py:59-63 through gym.make):
    accepted with the same names and defaults.  `render_mode` / `is_headless` choose between a GL window and the EGL
    offscreen observer there; here every pixel comes from the CUDA ray-caster (`render()`), so they are stored, not acted
    on.  `width` / `height` must be 64: the only size the reference's make_env passes and the one the kernel is built for.
    `strict_capacity=True` makes `step` raise when a per-env contact / constraint-row buffer overflowed (the dropped
    entries are otherwise only counted, see `overflow()`; capacity = the reference model's nconmax=100, DAPG_assets.xml:4)."""

    def __init__(self, env_id: str, num_envs: int = 1, device: int | str | torch.device = 0, seed: int = 0,
                 variation_type=None, fp64: bool = False, auto_reset: bool = False, reset_on_done: bool = False,
                 env_id_offset: int = 0, frame_skip: int | None = None, model: Model | None = None,
                 render_mode=None, width: int = 64, height: int = 64, is_headless: bool = False, strict_capacity: bool = False):
        if env_id not in ENV_IDS:
            raise ValueError(f"unknown env id {env_id!r}; expected one of {sorted(ENV_IDS)}")
        if render_mode not in (None, "rgb_array", "human"):
            raise ValueError(f"unsupported render_mode {render_mode!r}")
        if (width, height) != (64, 64):
            raise ValueError("the pixel path renders 64x64 (helpers.py:63); other sizes are not built")
        if not torch.cuda.is_available():
            raise RuntimeError("AdroitVecEnv needs a CUDA device (no CPU fallback)")
        self.env_id = env_id
        self.render_mode, self.width, self.height, self.is_headless = render_mode, width, height, is_headless
        self.strict_capacity = strict_capacity
        self.device = torch.device(device if not isinstance(device, int) else f"cuda:{device}")
        self.num_envs = int(num_envs)
        self.model = model if model is not None else load_compiled(ENV_IDS[env_id])
        # variation_type: only HammerEnvV0 acts on it (hammer_v0.py:110-129); door/pen/relocate store and ignore it
        # (door_v0.py:29, pen_v0.py:32, relocate_v0.py:22)
        self.variation_type = variation_type
        if variation_type is not None and env_id == "hammer-v0":
            self.model = self.model.with_variation(variation_type)
        if frame_skip is not None:        # a private copy: a caller-supplied `model` is never edited in place
            self.model = Model(task=self.model.task, names=self.model.names, A=dict(self.model.A))
            self.model.A["sizes"] = self.model.A["sizes"].copy()
            self.model.A["sizes"][9] = frame_skip
        self.abi = cuda_abi()
        flags = (_cabi.F64 if fp64 else 0) | (_cabi.AUTO_RESET if auto_reset else 0) | (_cabi.RESET_ON_DONE if reset_on_done else 0)
        blob = self.model.to_blob()
        h = C.c_void_p()
        self.abi.check(self.abi.create(blob, len(blob), self.num_envs, self.device.index or 0, seed, env_id_offset, flags, C.byref(h)))
        self._h = h
        sz = np.zeros(8, dtype=np.int32)
        self.abi.sizes(h, sz.ctypes.data)
        self.nq = self.nv = int(sz[0])
        self.nu, self.obs_dim, self.npar = int(sz[1]), int(sz[2]), int(sz[3])
        self.horizon, self.frame_skip, self.nbody, self.nsite = int(sz[4]), int(sz[5]), int(sz[6]), int(sz[7])
        self.max_episode_steps = self.horizon
        self.dt = 0.002 * self.frame_skip
        self.action_space = Box(-1.0, 1.0, (self.nu,))                      # hammer_v0.py:51-52
        self.observation_space = Box(-np.inf, np.inf, (self.obs_dim,))
        n = self.num_envs
        self._obs = torch.zeros((n, self.obs_dim), dtype=torch.float32, device=self.device)
        self._rew = torch.zeros(n, dtype=torch.float32, device=self.device)
        self._done = torch.zeros(n, dtype=torch.uint8, device=self.device)
        self._goal = torch.zeros(n, dtype=torch.uint8, device=self.device)

    # ------------------------------------------------------------------ lifecycle
    def close(self):
        if getattr(self, "_h", None):
            self.abi.destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    # ------------------------------------------------------------------ gym-style API (batched)
    def reset(self, mask: torch.Tensor | None = None):
        mp = None
        if mask is not None:
            mask = mask.to(device=self.device, dtype=torch.uint8).contiguous()
            mp = mask.data_ptr()
        self.abi.check(self.abi.reset(self._h, mp, self._obs.data_ptr(), self._stream()))
        return self._obs.clone(), {}

    def step(self, action: torch.Tensor):
        a = torch.as_tensor(action, dtype=torch.float32, device=self.device).reshape(self.num_envs, self.nu).contiguous()
        self.abi.check(self.abi.step(self._h, a.data_ptr(), self._obs.data_ptr(), self._rew.data_ptr(), self._done.data_ptr(),
                                     self._goal.data_ptr(), self._stream()))
        if self.strict_capacity:
            self.check_capacity()
        return self._obs.clone(), self._rew.clone(), self._done.bool(), {"goal_achieved": self._goal.bool()}

    def check_capacity(self):
        """Raise if any env ever had more contacts / constraint rows than the kernels keep (results would differ from the reference's)."""
        o = self.overflow()
        if o.sum():
            raise RuntimeError(f"capacity overflow: {int(o[0])} contacts and {int(o[1])} constraint rows were dropped")

    def forward(self):
        """sim.forward() (hammer_v0.py:131): mj_forward on the current state; refreshes the frames get_obs / get_env_state read."""
        self.abi.check(self.abi.forward(self._h, self._stream()))

    def step_host(self, action: np.ndarray, obs: np.ndarray, reward: np.ndarray, done: np.ndarray, goal: np.ndarray):
        """Reference-facing call with HOST buffers (pinned for full speed): H2D action, step, D2H results, synchronised."""
        self.abi.check(self.abi.step_host(self._h, action.ctypes.data, obs.ctypes.data, reward.ctypes.data, done.ctypes.data,
                                          goal.ctypes.data, self._stream()))

    def render(self, out: torch.Tensor | None = None):
        """64x64 RGB pixel observation of every env, u8 [N, 64, 64, 3] (hammer_v0.py:177-181 -> HeadlessObserver.render,
        headless_observer.py:34-52, then PlaNet's 128->64 resize); ray-cast on the device, see csrc/render.cuh."""
        if out is None:
            out = torch.empty((self.num_envs, 64, 64, 3), dtype=torch.uint8, device=self.device)
        self.abi.check(self.abi.render(self._h, out.data_ptr(), self._stream()))
        return out

    def get_obs(self):
        self.abi.check(self.abi.get_obs(self._h, self._obs.data_ptr(), self._stream()))
        return self._obs.clone()

    # ------------------------------------------------------------------ state
    def _raw_state(self):
        n = self.num_envs
        q, v, p = np.zeros((n, self.nv)), np.zeros((n, self.nv)), np.zeros((n, self.npar))
        self.abi.check(self.abi.get_state(self._h, q.ctypes.data, v.ctypes.data, p.ctypes.data, self._stream()))
        return q, v, p

    def _set_raw_state(self, q=None, v=None, p=None):
        c = lambda a, w: None if a is None else np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(self.num_envs, w))
        q, v, p = c(q, self.nv), c(v, self.nv), c(p, self.npar)
        g = lambda a: None if a is None else a.ctypes.data
        self.abi.check(self.abi.set_state(self._h, g(q), g(v), g(p), self._stream()))

    def get_env_state(self):
        """The reference's state dicts, batched: every value a float64 tensor [N, ...] (host).  Like the reference, no
        forward is run: `target_pos` / `obj_pos` / `palm_pos` are data.site_xpos / body_xpos as the last mj_forward left
        them, one substep behind qpos after a step (hammer_v0.py:134-142, relocate_v0.py:105-116)."""
        q, v, p = self._raw_state()
        task = ENV_IDS[self.env_id]
        T = lambda a: torch.from_numpy(np.ascontiguousarray(a))
        if task == "hammer":     # hammer_v0.py:134-142
            _, _, sx = self.frames()
            return dict(qpos=T(q), qvel=T(v), board_pos=T(p[:, :3]), target_pos=T(sx[:, self.model.name2id("site", "S_target")]))
        if task == "door":       # door_v0.py:121-128
            return dict(qpos=T(q), qvel=T(v), door_body_pos=T(p[:, :3]))
        if task == "pen":        # pen_v0.py:134-141
            return dict(qpos=T(q), qvel=T(v), desired_orien=T(p[:, :4]))
        bx, _, sx = self.frames()  # relocate_v0.py:105-116
        m = self.model
        return dict(hand_qpos=T(q[:, :30]), obj_pos=T(bx[:, m.name2id("body", "Object")]),
                    target_pos=T(sx[:, m.name2id("site", "target")]), palm_pos=T(sx[:, m.name2id("site", "S_grasp")]),
                    qpos=T(q), qvel=T(v))

    def set_env_state(self, state_dict):
        """set_state(qpos, qvel) + the model fields of the dict + sim.forward(); values may be tensors or arrays [N, ...]."""
        A = lambda x: x.detach().cpu().numpy() if isinstance(x, torch.Tensor) else np.asarray(x)
        task = ENV_IDS[self.env_id]
        _, _, p = self._raw_state()
        if task == "hammer":     # hammer_v0.py:144-153
            p[:, :3] = A(state_dict["board_pos"])
        elif task == "door":     # door_v0.py:130-138
            p[:, :3] = A(state_dict["door_body_pos"])
        elif task == "pen":      # pen_v0.py:143-152
            p[:, :4] = A(state_dict["desired_orien"])
        else:                    # relocate_v0.py:118-129 (writes the world-frame obj_pos into model.body_pos)
            p[:, :3] = A(state_dict["obj_pos"])
            p[:, 3:6] = A(state_dict["target_pos"])
        self._set_raw_state(A(state_dict["qpos"]), A(state_dict["qvel"]), p)

    # ------------------------------------------------------------------ success / statistics
    def evaluate_success(self, paths):
        return evaluate_success(self.env_id, paths)

    def reduce_episode(self, flush_running=False, all_reduce=True):
        """[sum of episode returns, #episodes with any goal, #episodes with >thr goal steps, #episodes] over all envs
        (and over all ranks with one NCCL all-reduce when torch.distributed is initialised; eval.py:69-97 semantics)."""
        out = np.zeros(4)
        self.abi.check(self.abi.reduce_episode(self._h, int(flush_running), out.ctypes.data, self._stream()))
        if self.strict_capacity:
            self.check_capacity()
        if all_reduce:
            from .sharding import allreduce_stats
            out = allreduce_stats(out, self.device)
        return out

    def reduce_episode_noalloc(self, out4: torch.Tensor, flush_running=False):
        """The statistics of reduce_episode written into the caller's device tensor `out4` (4 x float64) by
        adroit_reduce_episode_noalloc: no allocation, no host copy, no synchronisation, no all-reduce (usable under
        CUDA graph capture; limit ADROIT_NOALLOC_MAX_ENVS = 65536 envs per handle, calls must not overlap on the GPU)."""
        if out4.dtype != torch.float64 or out4.numel() != 4 or not out4.is_cuda or not out4.is_contiguous():
            raise ValueError("out4 must be a contiguous CUDA tensor of 4 float64")
        self.abi.check(self.abi.reduce_episode_noalloc(self._h, int(flush_running), out4.data_ptr(), self._stream()))
        return out4

    # ------------------------------------------------------------------ introspection (parity tests)
    def frames(self):
        n = self.num_envs
        bx, bq, sx = np.zeros((n, self.nbody, 3)), np.zeros((n, self.nbody, 4)), np.zeros((n, self.nsite, 3))
        self.abi.check(self.abi.get_frames(self._h, bx.ctypes.data, bq.ctypes.data, sx.ctypes.data, self._stream()))
        return bx, bq, sx

    def solver(self):
        n = self.num_envs
        qacc, ncon, nefc, touch = np.zeros((n, self.nv)), np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n)
        self.abi.check(self.abi.get_solver(self._h, qacc.ctypes.data, ncon.ctypes.data, nefc.ctypes.data, touch.ctypes.data, self._stream()))
        return qacc, ncon, nefc, touch

    def set_elapsed_steps(self, t):
        """Per-env episode step counters (TimeLimit._elapsed_steps); with auto_reset an env resets when its counter reaches the horizon."""
        t = np.ascontiguousarray(np.asarray(t, dtype=np.int32).reshape(self.num_envs))
        self.abi.check(self.abi.set_elapsed_steps(self._h, t.ctypes.data, self._stream()))

    def get_warmstart(self):
        w = np.zeros((self.num_envs, self.nv))
        self.abi.check(self.abi.get_warmstart(self._h, w.ctypes.data, self._stream()))
        return w

    def set_warmstart(self, w):
        w = np.ascontiguousarray(np.asarray(w, dtype=np.float64).reshape(self.num_envs, self.nv))
        self.abi.check(self.abi.set_warmstart(self._h, w.ctypes.data, self._stream()))

    def overflow(self):
        o = np.zeros(2, dtype=np.int64)
        self.abi.check(self.abi.overflow(self._h, o.ctypes.data, self._stream()))
        return o

    def profile(self, enable=True):
        self.abi.check(self.abi.profile(self._h, int(enable)))

    def phase_times(self):
        """(ms[4], launches[4]) of the smooth / collide / solve / finish kernels since the last call (needs profile(True))."""
        ms, n = np.zeros(4), np.zeros(4, dtype=np.int64)
        self.abi.check(self.abi.phase_times(self._h, ms.ctypes.data, n.ctypes.data, self._stream()))
        return ms, n

    def launch_count(self):
        return int(self.abi.launch_count(self._h))

    def debug_scratch(self, offset, count):
        out = np.zeros((self.num_envs, count))
        self.abi.check(self.abi.lib.adroit_debug_scratch(self._h, offset, count, out.ctypes.data, self._stream()))
        return out

    def debug_layout(self):
        v = np.zeros(64, dtype=np.int32)
        n = self.abi.lib.adroit_debug_layout(v.ctypes.data, 64)
        keys = ["NVP", "NBP", "NGP", "NSP", "MAXCON", "CROWS", "NTRI", "NPARP", "NTF", "G_TOTAL", "GI_TOTAL", "G_GPOS", "G_GMAT", "G_SPOS",
                "G_CON", "G_JR", "G_MINV", "G_DBG_M", "G_DBG_BIAS", "G_DBG_QFS", "G_DBG_QAS", "G_DBG_QACC", "G_DBG_QFC", "G_DBG_XPOS",
                "G_DBG_XQUAT", "G_DBG_EF", "GI_CON", "SMEM_SMOOTH", "SMEM_COLLIDE", "SMEM_SOLVE_S", "SMEM_SOLVE", "SMEM_FINISH", "SMEM_RESET",
                "CTX_T", "CTX_I"]
        return dict(zip(keys, map(int, v[:n])))


def make(env_id: str, num_envs: int = 1, **kw) -> AdroitVecEnv:
    """gym.make-style constructor for the registered ids (mj_envs_vision/__init__.py:4-29); takes the reference's
    keyword arguments, so helpers.py:59-63's call `gym.make(id, is_headless=..., variation_type=..., render_mode='rgb_array',
    width=64, height=64)` works verbatim."""
    return AdroitVecEnv(env_id, num_envs, **kw)
