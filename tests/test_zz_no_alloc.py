"""adroit_reduce_episode_noalloc -- the allocation-free variant of adroit_reduce_episode (include/adroit_b200.h): its partial
sums live in a file-scope `static __device__` array of the library instead of cudaMalloc'ed memory.  No counterpart in the
reference, so the checker here is the existing entry point, called on the same
handle.  Named test_zz_* so that pytest collects it last: a failure here must not hide another
test.  Every test builds its own handles, so the file passes alone in a fresh process as well as after the other files."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-9          # the return sum is the same fp64 terms added in a different order; the three counts must be equal


def _mods():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from mj_envs_b200.vec_env import AdroitVecEnv
    return torch, AdroitVecEnv


def _rollout(torch, env, rng, steps):
    for _ in range(steps):
        a = torch.from_numpy(rng.uniform(-1, 1, (env.num_envs, env.nu)).astype(np.float32))
        env.step(a)


def _noalloc(torch, env, flush, out=None):
    """The new entry point through the C ABI, into a caller-owned device buffer that starts as NaN."""
    if out is None:
        out = torch.full((4,), float("nan"), dtype=torch.float64, device=env.device)
    stream = torch.cuda.current_stream(env.device).cuda_stream
    rc = env.abi.reduce_episode_noalloc(env._h, int(flush), out.data_ptr(), stream)
    assert rc == 0, env.abi.last_error()
    torch.cuda.synchronize(env.device)
    return out.cpu().numpy()


def _same(got, want):
    assert np.all(np.isfinite(got)), got
    assert np.array_equal(got[1:], want[1:]), (got, want)
    assert abs(got[0] - want[0]) <= RTOL * max(1.0, abs(want[0])), (got, want)


def test_noalloc_reduce_matches_reduce_episode_on_two_calls_in_a_row():
    torch, AdroitVecEnv = _mods()
    rng = np.random.default_rng(11)
    # 2000 envs = 8 rows of the static scratch (the last one ragged); pen with reset-on-done finishes episodes early
    env = AdroitVecEnv("pen-v0", 2000, seed=3, auto_reset=True, reset_on_done=True)
    env.reset()
    env.set_elapsed_steps(rng.integers(0, env.horizon, env.num_envs))   # staggered: some envs reach the horizon in every step
    _rollout(torch, env, rng, 30)
    want1 = env.reduce_episode(flush_running=True, all_reduce=False)
    got1 = _noalloc(torch, env, True)
    assert want1[3] > env.num_envs          # finished episodes beyond the ones in flight: both terms of the sum are exercised
    _same(got1, want1)
    _same(_noalloc(torch, env, False), env.reduce_episode(flush_running=False, all_reduce=False))

    # second round on the same handle: the scratch still holds the first round's rows, the statistics have moved on
    _rollout(torch, env, rng, 12)
    want2 = env.reduce_episode(flush_running=True, all_reduce=False)
    bufs = [torch.full((4,), float("nan"), dtype=torch.float64, device=env.device) for _ in range(2)]
    torch.cuda.synchronize(env.device)
    free0 = torch.cuda.mem_get_info(env.device)[0]
    got2 = _noalloc(torch, env, True, bufs[0])
    got2b = _noalloc(torch, env, True, bufs[1])
    assert torch.cuda.mem_get_info(env.device)[0] == free0      # nothing was allocated on the device by the two calls
    assert want2[3] > want1[3]
    _same(got2, want2)
    assert np.array_equal(got2, got2b)                           # fixed summation order: bit-reproducible

    # a smaller fp64 handle right after a larger one: rows 1..7 of the scratch are stale and must not leak into the result
    small = AdroitVecEnv("door-v0", 40, seed=5, fp64=True, auto_reset=True)
    small.reset()
    _rollout(torch, small, rng, 5)
    _same(_noalloc(torch, small, True), small.reduce_episode(flush_running=True, all_reduce=False))
    # the wrapper of the front end is the same call
    out = torch.empty(4, dtype=torch.float64, device=small.device)
    small.reduce_episode_noalloc(out, flush_running=True)
    torch.cuda.synchronize(small.device)
    _same(out.cpu().numpy(), small.reduce_episode(flush_running=True, all_reduce=False))
    # and the first handle again, untouched by the other handle's call
    _same(_noalloc(torch, env, True), want2)
    small.close()
    env.close()


def test_noalloc_reduce_rejects_more_envs_than_the_static_scratch_holds():
    torch, AdroitVecEnv = _mods()
    import gc
    gc.collect()
    torch.cuda.empty_cache()
    limit = 65536                                                # ADROIT_NOALLOC_MAX_ENVS of include/adroit_b200.h
    hdr = open(__file__.rsplit("/tests/", 1)[0] + "/include/adroit_b200.h").read()
    assert f"#define ADROIT_NOALLOC_MAX_ENVS {limit}" in hdr
    at = AdroitVecEnv("door-v0", limit, seed=1)                  # exactly the limit: accepted (nothing stepped: all zeros)
    got = _noalloc(torch, at, False)
    assert np.array_equal(got, np.zeros(4))
    at.close()
    over = AdroitVecEnv("door-v0", limit + 1, seed=1)            # no reset, no step: only the context blocks are allocated
    out = torch.full((4,), -7.0, dtype=torch.float64, device=over.device)
    launches = over.abi.launch_count(over._h)
    rc = over.abi.reduce_episode_noalloc(over._h, 1, out.data_ptr(), torch.cuda.current_stream(over.device).cuda_stream)
    msg = over.abi.last_error().decode()
    torch.cuda.synchronize(over.device)
    assert rc != 0
    assert "ADROIT_NOALLOC_MAX_ENVS" in msg and str(limit + 1) in msg, msg
    assert over.abi.launch_count(over._h) == launches            # nothing was launched
    assert np.array_equal(out.cpu().numpy(), np.full(4, -7.0))   # the caller's buffer is untouched
    assert np.all(np.isfinite(over.reduce_episode(flush_running=True, all_reduce=False)))   # the allocating entry point still serves it
    with pytest.raises(RuntimeError, match="ADROIT_NOALLOC_MAX_ENVS"):
        over.reduce_episode_noalloc(out)
    with pytest.raises(ValueError):
        over.reduce_episode_noalloc(torch.zeros(4, dtype=torch.float32, device=over.device))
    over.close()
    assert C.sizeof(C.c_double) * 4 * (limit // 256) == 8192     # the stated size of the scratch: 8 KB
