"""Multi-GPU plumbing of the step path: envs are independent, so a job of `total` envs is cut into
contiguous per-rank ranges with NO data-path collective; the reset RNG is keyed on the global env id, so results do not
depend on the rank count.  The only exchange is the end-of-rollout all-reduce(sum) of the four episode statistics
(eval.py:69-97 / evaluate_success semantics): [sum of returns, #episodes with any goal, #successful episodes, #episodes]."""
from __future__ import annotations

import numpy as np


def shard_range(total_envs: int, rank: int, world: int) -> tuple[int, int]:
    """(env_id_offset, num_envs) of `rank`: contiguous, sizes differ by at most one, union = range(total_envs)."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    base, rem = divmod(total_envs, world)
    n = base + (1 if rank < rem else 0)
    off = rank * base + min(rank, rem)
    return off, n


def allreduce_stats(stats: np.ndarray, device=None) -> np.ndarray:
    """Sum the per-rank statistics over all ranks (one NCCL all-reduce of 4 fp64 over NVLink; gloo on CPU in the tests)."""
    import torch
    import torch.distributed as dist
    t = torch.from_numpy(np.ascontiguousarray(stats, dtype=np.float64))
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        if dist.get_backend() == "nccl":
            t = t.to(device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        t = t.cpu()
    return t.numpy()


def success_percentage(stats: np.ndarray) -> float:
    """evaluate_success (hammer_v0.py:167-175) over every episode of every rank."""
    return 100.0 * float(stats[2]) / max(float(stats[3]), 1.0)
