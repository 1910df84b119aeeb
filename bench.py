#!/usr/bin/env python
"""bench.py — Adroit env-steps/s of the batched step path on N B200s (one process per GPU), next to the reference
path's CPU restatement on the box's host cores.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload door-v0:4096] [--impl reference]

One "step" is one env.step() over the whole batch (action map -> frame_skip x mj_step -> obs/reward/success, with
auto-reset at the horizon), i.e. `envs` env-steps.  The default workload is BASELINE.json configs[1]: door-v0, 4096
envs per GPU, state observations, U(-1,1) actions.  Prints ONE JSON line (rank 0).
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
# algorithmic bytes per env-step: fp32 words read (action, qpos, qvel, warmstart, params) + written
# (qpos, qvel, warmstart, obs, reward, done+goal) x 4 B
ALGO_BYTES = {"hammer-v0": 1092, "door-v0": 1008, "pen-v0": 1020, "relocate-v0": 1168}


def parse_workload(s):
    env_id, _, n = s.partition(":")
    return env_id, int(n or 4096)


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


def cpu_reference_run(env_id, sample_envs, steps, warmup, seed=0):
    """The reference path's CPU restatement (oracle/liboracle.so: fp64, OpenMP over envs) on the host cores.
    mujoco-py itself is not installable on this box, hence kind="port"."""
    from oracle.oracle import OracleSim
    task = env_id.split("-")[0]
    cores = len(os.sched_getaffinity(0))
    # all the host threads this process may use: torchrun exports OMP_NUM_THREADS=1 for N>1, which would silently turn
    # the CPU arm into a single-thread run, so it is overridden here and the thread count the runtime reports is recorded
    os.environ["OMP_NUM_THREADS"] = str(cores)
    sim = OracleSim(task, sample_envs, seed=seed, flags=2)
    try:
        import ctypes
        gomp = ctypes.CDLL("libgomp.so.1")
        gomp.omp_set_num_threads(cores)
        cores = int(gomp.omp_get_max_threads())
    except OSError:
        pass
    sim.reset()
    rng = np.random.default_rng(seed)
    acts = rng.uniform(-1, 1, (8, sample_envs, sim.nu)).astype(np.float32)
    for i in range(warmup):
        sim.step(acts[i % 8])
    t0 = time.perf_counter()
    for i in range(steps):
        sim.step(acts[i % 8])
    dt = time.perf_counter() - t0
    return sample_envs * steps / dt, dt, cores


def run_reference(args, env_id, n_envs, rank, world):
    if rank != 0:
        return
    sample = min(n_envs, 1024)
    # keep the whole run within a few minutes: calibrate, then bound the step count
    v, dt, cores = cpu_reference_run(env_id, sample, 2, 1)
    per_step = dt / 2
    steps = max(1, min(args.steps, int(120.0 / max(per_step, 1e-6))))
    warm = max(1, min(args.warmup, 3))
    v, dt, cores = cpu_reference_run(env_id, sample, steps, warm)
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warm,
        "ms_per_step": 1e3 * dt / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": {"workload": f"{env_id} x{sample} envs (bounded sample of {n_envs}/GPU), state obs, U(-1,1) actions, "
                                        "auto-reset at the horizon"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{sample} envs x {steps} env.step() calls, OpenMP over envs; mujoco-py unavailable on this box"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="door-v0:4096", help="env_id:envs_per_gpu (default = BASELINE.json configs[1])")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--pixels", action="store_true", help="also produce the 64x64x3 u8 pixel observation every step (BASELINE.json configs[4])")
    ap.add_argument("--no-flush", action="store_true", help="do not flush L2 between timed steps")
    ap.add_argument("--also", default="hammer-v0:8192", help="comma list of extra workloads reported under 'also' (N=1 only; '' = none)")
    args = ap.parse_args()
    env_id, n_envs = parse_workload(args.workload)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, env_id, n_envs, rank, world)
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

    def measure(env_id, n_envs, steps, warmup, with_e2e):
        env = AdroitVecEnv(env_id, n_envs, device=local, seed=0, auto_reset=True, env_id_offset=rank * n_envs)
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

        rgb = torch.empty((n_envs, 64, 64, 3), dtype=torch.uint8, device=dev) if args.pixels else None

        def raw_step(i):
            env.abi.check(env.abi.step(h, acts[i % nact].data_ptr(), obs.data_ptr(), rew.data_ptr(), done.data_ptr(), goal.data_ptr(),
                                       stream.cuda_stream))
            if rgb is not None:
                env.abi.check(env.abi.render(h, rgb.data_ptr(), stream.cuda_stream))
        env.reset()
        for i in range(warmup):
            raw_step(i)
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
            raw_step(warmup + i)
            ev[i][1].record(stream)
        barrier()
        wall = time.perf_counter() - t_wall0
        launches = env.launch_count() - l0
        clocks = sampler.stop()
        ms = np.array([a.elapsed_time(b) for a, b in ev])
        total_ms = float(ms.sum())
        if world > 1:
            t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            total_ms = float(t.item())
        stats = env.reduce_episode(flush_running=True)   # one all-reduce of 4 numbers (eval.py:69-97 semantics)
        overflow = env.overflow()
        res = {"value": world * n_envs * steps / (total_ms * 1e-3), "ms_per_step": total_ms / steps, "launches": launches,
               "kernel_ms": float(ms.mean()), "clocks": clocks, "wall_s": wall, "stats": stats.tolist(), "overflow": overflow.tolist(),
               "nu": env.nu, "obs_dim": env.obs_dim}
        # per-phase device time of the same workload (CUDA events around every phase-kernel launch, on the launching stream)
        # (every measurement section restarts from reset + the same warm-up, so that all of them time the same steps of
        # the episodes: the cost of a step depends on where in the episode it is)
        env.reset()
        for i in range(warmup):
            raw_step(i)
        env.profile(True)
        psteps = min(steps, 40)
        for i in range(psteps):
            if flush is not None:
                flush.fill_(i & 0xFF)
            raw_step(warmup + i)
        pms, pn = env.phase_times()
        env.profile(False)
        res["phase_ms_per_launch"] = (pms / (psteps * env.frame_skip)).tolist()   # the masked reset pass (a few us) is folded in
        res["phase_share"] = (pms / pms.sum()).tolist()
        res["frame_skip"] = env.frame_skip
        if with_e2e:
            # end to end through the reference-facing call: pinned HOST buffers, H2D action + D2H obs/reward/done/goal every step
            a_h = [acts[j].cpu().pin_memory() for j in range(nact)]     # the same action sequence as the device-resident run
            o_h = torch.empty((n_envs, env.obs_dim), dtype=torch.float32).pin_memory()
            r_h = torch.empty(n_envs, dtype=torch.float32).pin_memory()
            d_h = torch.empty(n_envs, dtype=torch.uint8).pin_memory()
            g_h = torch.empty(n_envs, dtype=torch.uint8).pin_memory()

            rgb_h = torch.empty((n_envs, 64, 64, 3), dtype=torch.uint8).pin_memory() if args.pixels else None

            def host_step(i):
                if rgb is not None:   # pixels of the state this action is chosen from, then the step (both synchronise)
                    env.abi.check(env.abi.render(h, rgb.data_ptr(), stream.cuda_stream))
                    rgb_h.copy_(rgb, non_blocking=True)
                env.abi.check(env.abi.step_host(h, a_h[i % nact].data_ptr(), o_h.data_ptr(), r_h.data_ptr(), d_h.data_ptr(), g_h.data_ptr(),
                                                stream.cuda_stream))
            env.reset()
            for i in range(max(3, warmup)):
                host_step(i)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            e0.record(stream)
            acc = 0.0
            for i in range(steps):
                host_step(warmup + i)
                acc += float(r_h[0])            # the host consumes the step's result
            e1.record(stream)
            barrier()
            e2e_ms = e0.elapsed_time(e1)
            if world > 1:
                t = torch.tensor([e2e_ms], device=dev, dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                e2e_ms = float(t.item())
            res["e2e"] = {"value": world * n_envs * steps / (e2e_ms * 1e-3), "unit": UNIT,
                          "h2d_bytes_per_step": n_envs * env.nu * 4, "d2h_bytes_per_step": n_envs * (env.obs_dim * 4 + 4 + 1 + 1 + (64 * 64 * 3 if args.pixels else 0))}
        env.close()
        return res

    r = measure(env_id, n_envs, args.steps, args.warmup, True)
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    # dominant kernel: one launch advances n_envs by one physics substep = 1/frame_skip of an env-step
    names = ["smooth_kernel", "collide_kernel", "solve_kernel", "finish_kernel"]
    dom = int(np.argmax(r["phase_ms_per_launch"]))
    bytes_per_launch = ALGO_BYTES[env_id] * n_envs / r["frame_skip"]
    pix = 64 * 64 * 3 if args.pixels else 0   # algorithmic bytes of the pixel observation, per env-step
    achieved = bytes_per_launch / (r["phase_ms_per_launch"][dom] * 1e-3) / 1e9
    step_achieved = (ALGO_BYTES[env_id] + pix) * n_envs / (r["kernel_ms"] * 1e-3) / 1e9
    line = {
        "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": f"{env_id} x{n_envs} envs per GPU, state obs, U(-1,1) actions, auto-reset at the horizon "
                               "(BASELINE.json configs[1])" if args.workload == "door-v0:4096" else
                               f"{env_id} x{n_envs} envs per GPU, state obs{' + 64x64x3 u8 pixel obs (ray-cast)' if args.pixels else ''}, "
                               "U(-1,1) actions, auto-reset at the horizon",
                   "envs_per_gpu": n_envs, "l2": "flushed between timed steps (256 MiB fill)" if not args.no_flush else "not flushed",
                   "timing": "sum of per-step CUDA-event intervals on the launching stream, max over ranks"},
        "e2e": r["e2e"], "gpu_launches": r["launches"], "clocks": r["clocks"],
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                     "peak_source": peak_src, "algorithmic_bytes_per_env_step": ALGO_BYTES[env_id],
                     "kernel": names[dom], "kernel_ms_per_launch": r["phase_ms_per_launch"][dom], "algorithmic_bytes_per_launch": bytes_per_launch,
                     "kernel_share_of_step": r["phase_share"][dom], "phase_ms_per_launch": dict(zip(names, r["phase_ms_per_launch"])),
                     "whole_step_achieved": step_achieved, "whole_step_frac": step_achieved / peak,
                     "note": "the path is FP32-issue/latency bound, not HBM bound (DESIGN.md §5); reported as mandated"},
        "episode_stats": {"sum_return": r["stats"][0], "episodes_any_goal": r["stats"][1], "episodes_success": r["stats"][2],
                          "episodes": r["stats"][3]},
        "overflow": r["overflow"],
    }
    traffic_path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(traffic_path):
        try:
            # measured DRAM bytes of one launch of the dominant kernel (ncu --set full, profiles/), scaled from the
            # captured sub-batch launch to this launch's env count -- per launch, like `achieved`
            t = json.load(open(traffic_path)).get(env_id)
            if t and t.get("kernel") == names[dom]:
                line["roofline"]["traffic"] = t["dram_bytes_per_env_launch"] * n_envs
                line["roofline"]["traffic_source"] = t["source"]
        except Exception:
            pass
    if rank == 0 and world == 1 and args.also:
        line["also"] = {}
        for w in args.also.split(","):
            e2, n2 = parse_workload(w)
            r2 = measure(e2, n2, max(10, args.steps // 4), max(3, args.warmup // 4), False)
            line["also"][w] = {"value": r2["value"], "ms_per_step": r2["ms_per_step"], "overflow": r2["overflow"]}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        sample = min(n_envs, 1024)
        v, dt, cores = cpu_reference_run(env_id, sample, 4, 1)
        k = max(2, min(20000, int(12.0 / max(dt / 4, 1e-6))))        # ~12 s of CPU work
        v, dt, cores = cpu_reference_run(env_id, sample, k, 1)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": f"{env_id} x{sample} envs x {k} env.step() calls ({dt:.1f} s), fp64 oracle, OpenMP over envs; "
                                          "mujoco-py unavailable on this box"}
    elif rank == 0:
        line["cpu_baseline"] = None
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
