#!/usr/bin/env python
"""bench.py — Adroit env-steps/s of the batched step path on N B200s (one process per GPU), next to the reference
path's CPU restatement on the box's host cores.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload hammer-v0:8192] [--impl reference]

One "step" is one env.step() over the whole batch (action map -> frame_skip x mj_step -> obs/reward/success, with
auto-reset at the horizon), i.e. `envs` env-steps.  The default workload is the configuration BASELINE.json's target
is quoted on: hammer-v0, state observations, 8192 envs per GPU, U(-1,1) actions; the other BASELINE configs (door x4096
at this fork's frame_skip 1 and upstream's 5, pen x16384 with and without reset-on-drop, relocate x8192 per GPU =
configs[3] at 8 GPUs, hammer x8192 with pixels) are reported under "also", each with its own e2e and roofline.

Steady state: the cost of a step depends on where in the episode an env is (contacts build up), so the batch is
DE-PHASED before anything is timed: the episode step counters are staggered uniformly over the horizon and one full
horizon of steps is run (every env then resets at its own time, forever).  A window of any length, anywhere, sees the same
mix of episode phases; `--warmup` steps are run on top of that.  Prints ONE JSON line (rank 0).
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Adroit env-steps/sec"
UNIT = "env-steps/s"
DEFAULT_WORKLOAD = "hammer-v0:8192"
DEFAULT_ALSO = "door-v0:4096,door-v0:4096:fs5,pen-v0:16384,pen-v0:16384:rod,relocate-v0:8192,hammer-v0:8192:pixels"
# algorithmic bytes per env-step: fp32 words read (action, qpos, qvel, warmstart, params) + written
# (qpos, qvel, warmstart, obs, reward, done+goal) x 4 B
ALGO_BYTES = {"hammer-v0": 1092, "door-v0": 1008, "pen-v0": 1020, "relocate-v0": 1168}
PHASES = ["smooth_kernel", "collide_kernel", "solve_kernel", "finish_kernel"]


def parse_workload(s):
    """env_id:envs_per_gpu[:opt+opt]; opts: fs5 (door at upstream's frame_skip 5, door_v0.py:9-10), rod (reset on done: pen
    drop, pen_v0.py:94-98), pixels (64x64x3 u8 image per step, BASELINE configs[4])."""
    parts = s.split(":")
    env_id, n = parts[0], int(parts[1]) if len(parts) > 1 and parts[1] else 4096
    opts = set(parts[2].split("+")) if len(parts) > 2 and parts[2] else set()
    bad = opts - {"fs5", "rod", "pixels"}
    if bad or env_id not in ALGO_BYTES:
        raise SystemExit(f"bench.py: bad workload {s!r}")
    return env_id, n, opts


def describe(env_id, n, opts, what="envs per GPU"):
    s = f"{env_id} x{n} {what}, state obs"
    if "pixels" in opts: s += " + 64x64x3 u8 pixel obs (ray-cast)"
    s += ", U(-1,1) actions, auto-reset at the horizon"
    if "rod" in opts: s += " and on done (pen drop)"
    if "fs5" in opts: s += ", frame_skip 5 (upstream; this fork runs door at 1)"
    return s + ", episode phases staggered uniformly (steady state)"


class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz, self.ok = [], set(), None, False
        self._stop = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception as e:  # pragma: no cover
            self.err = repr(e)
        self.th = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                try:
                    r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                except Exception:
                    r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.05)

    def start(self):
        if self.ok:
            self.th.start()

    def stop(self):
        self._stop.set()
        if self.ok:
            self.th.join(timeout=2)
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def stagger(n, horizon, offset=0):
    """Episode step counters spread uniformly over [0, horizon), a function of the GLOBAL env id."""
    return (((np.arange(n, dtype=np.int64) + offset) * 2654435761) % horizon).astype(np.int32)


# ------------------------------------------------------------------------------------------------ CPU arm
def _omp_all_cores():
    cores = len(os.sched_getaffinity(0))
    # all the host threads this process may use: torchrun exports OMP_NUM_THREADS=1 for N>1, which would silently turn
    # the CPU arm into a single-thread run, so it is overridden here and the thread count the runtime reports is recorded
    os.environ["OMP_NUM_THREADS"] = str(cores)
    try:
        import ctypes
        gomp = ctypes.CDLL("libgomp.so.1")
        gomp.omp_set_num_threads(cores)
        cores = int(gomp.omp_get_max_threads())
    except OSError:
        pass
    return cores


class CpuArm:
    """The reference path's CPU restatement (oracle/liboracle.so: fp64, OpenMP over envs) on the host cores.
    mujoco-py itself is not installable on this box, hence kind="port"."""

    def __init__(self, env_id, opts, envs, seed=0, threads=None):
        from oracle.oracle import OracleSim
        from mj_envs_b200.mjcf import load_compiled, Model
        task = env_id.split("-")[0]
        self.cores = _omp_all_cores() if threads is None else threads
        if threads is not None:
            import ctypes
            ctypes.CDLL("libgomp.so.1").omp_set_num_threads(threads)
        m = load_compiled(task)
        if "fs5" in opts:
            m = Model(task=m.task, names=m.names, A=dict(m.A)); m.A["sizes"] = m.A["sizes"].copy(); m.A["sizes"][9] = 5
        self.sim = OracleSim(m, envs, seed=seed, flags=2 | (4 if "rod" in opts else 0))
        self.envs = envs
        rng = np.random.default_rng(seed)
        self.acts = rng.uniform(-1, 1, (8, envs, self.sim.nu)).astype(np.float32)
        self.i = 0
        self.sim.reset()

    def dephase(self, budget_s):
        """Stagger + pre-roll (as the GPU arm), as much of one horizon as the time budget allows."""
        self.sim.set_elapsed_steps(stagger(self.envs, self.sim.horizon))
        t0 = time.perf_counter()
        k = 0
        while k < self.sim.horizon and time.perf_counter() - t0 < budget_s:
            self.step(); k += 1
        return k

    def step(self):
        self.sim.step(self.acts[self.i % 8]); self.i += 1

    def run(self, steps):
        t0 = time.perf_counter()
        for _ in range(steps):
            self.step()
        return time.perf_counter() - t0


def cpu_config1(steps=1000):
    """BASELINE.json configs[0]: hammer-v0, ONE env, state obs, 1000 random-action steps, reset every 200, one host thread."""
    arm = CpuArm("hammer-v0", set(), 1, seed=0, threads=1)
    arm.run(20)
    dt = arm.run(steps)
    _omp_all_cores()
    return {"value": steps / dt, "unit": UNIT, "cores": 1, "steps": steps,
            "config": "BASELINE configs[0]: hammer-v0 single env, state obs, 1000 U(-1,1)-action steps, auto-reset every 200, fp64 oracle on 1 host thread "
                      "(mujoco-py: unavailable on this box)"}


def run_reference(args, env_id, n_envs, opts, rank, world):
    if rank != 0:
        return
    # same env count as the GPU arm when the run then still ends within a few minutes, else a bounded sample of it
    probe = CpuArm(env_id, opts, 256)
    per_env_step = probe.run(2) / (2 * 256)
    del probe
    budget = 150.0
    sample = n_envs
    while sample > 256 and per_env_step * sample * (args.steps + args.warmup + 30) > budget:
        sample //= 2
    arm = CpuArm(env_id, opts, sample)
    pre = arm.dephase(max(5.0, budget - per_env_step * sample * (args.steps + args.warmup)))
    arm.run(args.warmup)
    dt = arm.run(args.steps)
    v = sample * args.steps / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": describe(env_id, n_envs, opts), "envs_per_gpu": n_envs,
                   "sample": f"{sample} of {n_envs} envs per step" + ("" if sample == n_envs else " (bounded: the CPU arm's throughput does not depend on the batch size beyond a few envs per thread)"),
                   "pre_roll_steps": pre},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": arm.cores, "kind": "port", "per_core": v / arm.cores,
                         "sample": f"{sample} envs x {args.steps} env.step() calls after {pre} de-phasing + {args.warmup} warm-up steps, fp64 oracle, OpenMP over envs; "
                                   "mujoco-py unavailable on this box"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ GPU arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, help="env_id:envs_per_gpu[:opts] (default = the config the north-star target is quoted on)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--pixels", action="store_true", help="same as the :pixels workload option")
    ap.add_argument("--no-flush", action="store_true", help="do not flush L2 between timed steps")
    ap.add_argument("--no-dephase", action="store_true", help="time from the start of the episodes (all envs in phase), as round 1 did")
    ap.add_argument("--also", default=None, help="comma list of extra workloads reported under 'also' ('' or 'none' = none); default: the other BASELINE configs at N=1, "
                                                 "relocate-v0:8192 (= configs[3] at 8 GPUs) at N>1")
    ap.add_argument("--e2e-steps", type=int, default=200, help="the end-to-end legs time max(steps, this) steps")
    args = ap.parse_args()
    env_id, n_envs, opts = parse_workload(args.workload)
    if args.pixels:
        opts.add("pixels")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, env_id, n_envs, opts, rank, world)
        return

    import torch
    import torch.distributed as dist
    from mj_envs_b200.vec_env import AdroitVecEnv

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the CUDA path has no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[local])
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    traffic = {}
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    except Exception:
        pass

    def measure(env_id, n_envs, opts, steps, warmup, e2e_steps):
        pixels = "pixels" in opts
        env = AdroitVecEnv(env_id, n_envs, device=local, seed=0, auto_reset=True, reset_on_done="rod" in opts,
                           env_id_offset=rank * n_envs, frame_skip=5 if "fs5" in opts else None)
        gen = torch.Generator(device=dev)
        gen.manual_seed(1234 + rank)
        nact = 16
        acts = torch.rand((nact, n_envs, env.nu), device=dev, generator=gen) * 2 - 1   # resident in HBM before timing
        flush = None if args.no_flush else torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        obs = torch.empty((n_envs, env.obs_dim), dtype=torch.float32, device=dev)
        rew = torch.empty(n_envs, dtype=torch.float32, device=dev)
        done = torch.empty(n_envs, dtype=torch.uint8, device=dev)
        goal = torch.empty(n_envs, dtype=torch.uint8, device=dev)
        stream = torch.cuda.current_stream(dev)
        h = env._h
        rgb = torch.empty((n_envs, 64, 64, 3), dtype=torch.uint8, device=dev) if pixels else None
        counter = [0]

        def raw_step():
            i = counter[0]; counter[0] += 1
            env.abi.check(env.abi.step(h, acts[i % nact].data_ptr(), obs.data_ptr(), rew.data_ptr(), done.data_ptr(), goal.data_ptr(),
                                       stream.cuda_stream))
            if rgb is not None:
                env.abi.check(env.abi.render(h, rgb.data_ptr(), stream.cuda_stream))

        env.reset()
        pre = 0
        if not args.no_dephase:
            env.set_elapsed_steps(stagger(n_envs, env.horizon, rank * n_envs))
            pre = env.horizon
            for _ in range(pre):                # one full horizon: every env has been reset at its own time
                raw_step()
        for _ in range(warmup):
            raw_step()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        barrier()
        sampler = ClockSampler(local)
        sampler.start()
        l0 = env.launch_count()
        t_wall0 = time.perf_counter()
        for i in range(steps):
            if flush is not None:
                flush.fill_(i & 0xFF)          # L2 flush (256 MiB > 126 MB L2), outside the timed events
            ev[i][0].record(stream)
            raw_step()
            ev[i][1].record(stream)
        barrier()
        wall = time.perf_counter() - t_wall0
        launches = env.launch_count() - l0
        clocks = sampler.stop()
        ms = np.array([a.elapsed_time(b) for a, b in ev])
        total_ms = max_over_ranks(float(ms.sum()))
        res = {"value": world * n_envs * steps / (total_ms * 1e-3), "ms_per_step": total_ms / steps, "launches": launches,
               "kernel_ms": float(ms.mean()), "clocks": clocks, "wall_s": wall, "pre_roll_steps": pre, "nu": env.nu, "obs_dim": env.obs_dim,
               "ms_per_step_min_max": [float(ms.min()), float(ms.max())]}
        # per-phase device time (CUDA events around every phase-kernel launch, on the launching stream); the batch stays in
        # its de-phased steady state, so these are the same kind of steps as the ones timed above
        env.profile(True)
        psteps = min(steps, 40)
        for i in range(psteps):
            if flush is not None:
                flush.fill_(i & 0xFF)
            raw_step()
        pms, pn = env.phase_times()
        env.profile(False)
        # per physics substep over the whole batch (the solve phase is one launch per size class: their sum)
        res["phase_ms_per_launch"] = (pms / (psteps * env.frame_skip)).tolist()
        res["phase_launches_per_step"] = (pn / psteps).tolist()
        res["phase_share"] = (pms / pms.sum()).tolist()
        res["frame_skip"] = env.frame_skip
        # ---- end to end, (1) through the C ABI's reference-facing call: pinned HOST buffers, H2D action + D2H results every step
        a_h = [acts[j].cpu().pin_memory() for j in range(nact)]
        o_h = torch.empty((n_envs, env.obs_dim), dtype=torch.float32).pin_memory()
        r_h = torch.empty(n_envs, dtype=torch.float32).pin_memory()
        d_h = torch.empty(n_envs, dtype=torch.uint8).pin_memory()
        g_h = torch.empty(n_envs, dtype=torch.uint8).pin_memory()
        rgb_h = torch.empty((n_envs, 64, 64, 3), dtype=torch.uint8).pin_memory() if pixels else None

        def host_step():
            i = counter[0]; counter[0] += 1
            if rgb is not None:   # pixels of the state this action is chosen from, then the step (both synchronise)
                env.abi.check(env.abi.render(h, rgb.data_ptr(), stream.cuda_stream))
                rgb_h.copy_(rgb, non_blocking=True)
            env.abi.check(env.abi.step_host(h, a_h[i % nact].data_ptr(), o_h.data_ptr(), r_h.data_ptr(), d_h.data_ptr(), g_h.data_ptr(),
                                            stream.cuda_stream))

        def api_step():           # (2) through the Python API users call: AdroitVecEnv.step(host tensor) -> tensors, results read on the host
            i = counter[0]; counter[0] += 1
            if rgb is not None:
                rgb_h.copy_(env.render(rgb), non_blocking=True)
            o, r, d, info = env.step(a_h[i % nact])
            o_h.copy_(o, non_blocking=True); r_h.copy_(r, non_blocking=True)
            d_h.copy_(d.to(torch.uint8), non_blocking=True); g_h.copy_(info["goal_achieved"].to(torch.uint8), non_blocking=True)
            torch.cuda.current_stream(dev).synchronize()

        n_e2e = max(steps, e2e_steps)
        bytes_out = n_envs * (env.obs_dim * 4 + 4 + 1 + 1 + (64 * 64 * 3 if pixels else 0))
        for key, fn in (("e2e", host_step), ("e2e_api", api_step)):
            for _ in range(max(3, warmup)):
                fn()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            e0.record(stream)
            acc = 0.0
            for _ in range(n_e2e):
                fn()
                acc += float(r_h[0])            # the host consumes the step's result
            e1.record(stream)
            barrier()
            e2e_ms = max_over_ranks(e0.elapsed_time(e1))
            res[key] = {"value": world * n_envs * n_e2e / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": n_envs * env.nu * 4,
                        "d2h_bytes_per_step": bytes_out, "steps": n_e2e,
                        "through": "adroit_step_host (C ABI, pinned host buffers)" if key == "e2e" else "AdroitVecEnv.step (Python API) + host reads"}
        stats = env.reduce_episode(flush_running=True)   # one all-reduce of 4 numbers (eval.py:69-97 semantics)
        res["stats"] = stats.tolist()
        res["overflow"] = env.overflow().tolist()
        env.close()
        # ---- roofline of the dominant kernel (HBM, on algorithmic bytes, as mandated)
        dom = int(np.argmax(np.array(res["phase_share"])))
        bytes_per_launch = ALGO_BYTES[env_id] * n_envs / res["frame_skip"]
        pix = 64 * 64 * 3 if pixels else 0
        achieved = bytes_per_launch / (res["phase_ms_per_launch"][dom] * 1e-3) / 1e9
        step_achieved = (ALGO_BYTES[env_id] + pix) * n_envs / (res["kernel_ms"] * 1e-3) / 1e9
        roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "peak_source": peak_src, "algorithmic_bytes_per_env_step": ALGO_BYTES[env_id],
                "kernel": PHASES[dom], "kernel_ms_per_launch": res["phase_ms_per_launch"][dom], "algorithmic_bytes_per_launch": bytes_per_launch,
                "kernel_share_of_step": res["phase_share"][dom], "phase_ms_per_launch": dict(zip(PHASES, res["phase_ms_per_launch"])),
                "whole_step_achieved": step_achieved, "whole_step_frac": step_achieved / peak,
                "note": "the path is FP32-issue/latency bound, not HBM bound (DESIGN.md §5); reported as mandated"}
        t = traffic.get(env_id)
        if t and t.get("kernel") == PHASES[dom]:
            # measured DRAM bytes of one launch of the dominant kernel (ncu --set full, profiles/), scaled from the
            # captured launch to this launch's env count -- per launch, like `achieved`
            roof["traffic"] = t["dram_bytes_per_env_launch"] * n_envs
            roof["traffic_source"] = t["source"]
        res["roofline"] = roof
        return res

    r = measure(env_id, n_envs, opts, args.steps, args.warmup, args.e2e_steps)
    line = {
        "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": describe(env_id, n_envs, opts) + (" (the configuration BASELINE.json's north-star target is quoted on)" if args.workload == DEFAULT_WORKLOAD else ""),
                   "envs_per_gpu": n_envs, "pre_roll_steps": r["pre_roll_steps"],
                   "l2": "flushed between timed steps (256 MiB fill)" if not args.no_flush else "not flushed",
                   "timing": "sum of per-step CUDA-event intervals on the launching stream, max over ranks"},
        "e2e": r["e2e"], "e2e_api": r["e2e_api"], "gpu_launches": r["launches"], "clocks": r["clocks"],
        "roofline": r["roofline"],
        "ms_per_step_min_max": r["ms_per_step_min_max"],
        "episode_stats": {"sum_return": r["stats"][0], "episodes_any_goal": r["stats"][1], "episodes_success": r["stats"][2],
                          "episodes": r["stats"][3]},
        "overflow": r["overflow"],
    }
    also = args.also if args.also is not None else (DEFAULT_ALSO if world == 1 else "relocate-v0:8192")
    if also in ("none", "''", '""'):
        also = ""
    if also:
        line["also"] = {}
        for w in also.split(","):
            e2, n2, o2 = parse_workload(w)
            if (e2, n2, o2) == (env_id, n_envs, opts):
                continue
            r2 = measure(e2, n2, o2, max(10, args.steps // 2), max(3, args.warmup // 2), 100)
            entry = {"config": describe(e2, n2, o2), "value": r2["value"], "ms_per_step": r2["ms_per_step"], "e2e": r2["e2e"], "e2e_api": r2["e2e_api"],
                     "roofline": r2["roofline"], "gpu_launches": r2["launches"], "overflow": r2["overflow"],
                     "episode_stats": {"sum_return": r2["stats"][0], "episodes_any_goal": r2["stats"][1], "episodes_success": r2["stats"][2], "episodes": r2["stats"][3]}}
            if world > 1 and e2 == "relocate-v0":
                entry["config"] = describe(e2, n2 * world, o2, what=f"envs over {world} GPUs") + " -- BASELINE configs[3]; episode return / success all-reduced over NCCL"
            line["also"][w] = entry
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        sample = min(n_envs, 1024)
        arm = CpuArm(env_id, opts, sample)
        pre = arm.dephase(8.0)
        dt = arm.run(2)
        k = max(2, min(20000, int(12.0 / max(dt / 2, 1e-6))))        # ~12 s of CPU work
        dt = arm.run(k)
        v = sample * k / dt
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": arm.cores, "kind": "port", "per_core": v / arm.cores,
                                "sample": f"{env_id} x{sample} envs x {k} env.step() calls ({dt:.1f} s) after {pre} de-phasing steps, fp64 oracle, OpenMP over envs; "
                                          "mujoco-py unavailable on this box",
                                "config1": cpu_config1()}
    elif rank == 0:
        line["cpu_baseline"] = None
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
