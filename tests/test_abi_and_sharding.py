"""CPU-side checks: the C-ABI library loads and exports every symbol include/adroit_b200.h declares (no compute calls
without a GPU), the CUDA front-end fails loudly without a device, and the N>1 host logic (sharding + the end-of-rollout
all-reduce) gives rank-count-invariant results on a world_size-2 gloo group."""
import ctypes
import os
import re
import socket
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    hdr = open(os.path.join(ROOT, "include", "adroit_b200.h")).read()
    return sorted(set(re.findall(r"\b(adroit_[a-z_0-9]+)\s*\(", hdr)))


def test_cuda_library_exports_every_declared_symbol():
    from mj_envs_b200 import build
    lib = ctypes.CDLL(build.build())
    names = _declared()
    assert len(names) >= 17
    for n in names:
        assert hasattr(lib, n), n


def test_oracle_exports_the_same_abi():
    from oracle import oracle
    lib = ctypes.CDLL(oracle.build())
    for n in _declared():
        assert hasattr(lib, n.replace("adroit_", "oracle_", 1)), n


def test_front_end_fails_loudly_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from mj_envs_b200.vec_env import AdroitVecEnv
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        AdroitVecEnv("hammer-v0", 4)


def test_product_package_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "mj_envs_b200")
    # the strict form: no import / dlopen of anything under oracle/
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, re.M), f
                assert "liboracle" not in src, f


def test_shard_ranges_partition_the_job():
    from mj_envs_b200.sharding import shard_range
    for total, world in [(65536, 8), (10, 3), (7, 8), (4096, 1)]:
        ids = []
        for r in range(world):
            off, n = shard_range(total, r, world)
            ids += list(range(off, off + n))
        assert ids == list(range(total))
    with pytest.raises(ValueError):
        shard_range(8, 2, 2)


WORKER = r'''
import os, sys, numpy as np
sys.path.insert(0, sys.argv[1])
import torch.distributed as dist
from oracle.oracle import OracleSim
from mj_envs_b200.sharding import shard_range, allreduce_stats
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
total, T = 12, 30
off, n = shard_range(total, rank, world)
sim = OracleSim("pen", n, seed=13, env_id_offset=off, flags=2 | 4)
sim.reset()
rng = np.random.default_rng(0)
for t in range(T):
    a = rng.uniform(-1, 1, (total, sim.nu)).astype(np.float32)[off:off + n]
    sim.step(a)
out = allreduce_stats(sim.reduce_episode(flush=True))
if rank == 0:
    np.save(sys.argv[2], out)
dist.destroy_process_group()
'''


def test_two_rank_gloo_reduction_equals_single_process(tmp_path):
    from oracle.oracle import OracleSim
    total, T = 12, 30
    sim = OracleSim("pen", total, seed=13, flags=2 | 4)
    sim.reset()
    rng = np.random.default_rng(0)
    for t in range(T):
        sim.step(rng.uniform(-1, 1, (total, sim.nu)).astype(np.float32))
    want = sim.reduce_episode(flush=True)
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    out = tmp_path / "out.npy"
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    env = dict(os.environ, OMP_NUM_THREADS="2")
    subprocess.check_call([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                           "--master-port", str(port), str(script), ROOT, str(out)], env=env, timeout=300)
    got = np.load(out)
    assert got[3] == want[3] >= total
    assert got[1] == want[1] and got[2] == want[2]
    assert abs(got[0] - want[0]) < 1e-9 * max(1.0, abs(want[0]))
