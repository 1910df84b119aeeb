"""Where does the end-to-end (host-buffer) step time go?  door-v0 x 4096."""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np, torch
from mj_envs_b200.vec_env import AdroitVecEnv
env_id, n = (sys.argv[1] if len(sys.argv) > 1 else "door-v0"), int(sys.argv[2]) if len(sys.argv) > 2 else 4096
env = AdroitVecEnv(env_id, n, device=0, seed=0, auto_reset=True)
dev = torch.device("cuda:0")
env.reset()
st = torch.cuda.current_stream(dev)
acts = torch.rand((n, env.nu), device=dev) * 2 - 1
obs = torch.empty((n, env.obs_dim), device=dev); rew = torch.empty(n, device=dev)
done = torch.empty(n, dtype=torch.uint8, device=dev); goal = torch.empty(n, dtype=torch.uint8, device=dev)
h = env._h
def dev_step():
    env.abi.check(env.abi.step(h, acts.data_ptr(), obs.data_ptr(), rew.data_ptr(), done.data_ptr(), goal.data_ptr(), st.cuda_stream))
for _ in range(20): dev_step()
torch.cuda.synchronize()
K = 100
# (1) back-to-back device steps, no sync: device-bound throughput + CPU enqueue time
t0 = time.perf_counter()
for _ in range(K): dev_step()
t_enq = time.perf_counter() - t0
torch.cuda.synchronize()
t_all = time.perf_counter() - t0
print(f"device steps back to back: enqueue {1e3*t_enq/K:.3f} ms/step, total {1e3*t_all/K:.3f} ms/step")
# (2) device step + sync each step
t0 = time.perf_counter()
for _ in range(K):
    dev_step(); torch.cuda.synchronize()
print(f"device step + sync: {1e3*(time.perf_counter()-t0)/K:.3f} ms/step")
# (3) host step
a_h = torch.empty((n, env.nu)).pin_memory(); a_h.copy_(torch.rand((n, env.nu)) * 2 - 1)
o_h = torch.empty((n, env.obs_dim)).pin_memory(); r_h = torch.empty(n).pin_memory()
d_h = torch.empty(n, dtype=torch.uint8).pin_memory(); g_h = torch.empty(n, dtype=torch.uint8).pin_memory()
def host_step():
    env.abi.check(env.abi.step_host(h, a_h.data_ptr(), o_h.data_ptr(), r_h.data_ptr(), d_h.data_ptr(), g_h.data_ptr(), st.cuda_stream))
for _ in range(10): host_step()
t0 = time.perf_counter()
for _ in range(K): host_step()
print(f"host step: {1e3*(time.perf_counter()-t0)/K:.3f} ms/step")
# (4) copies alone
t0 = time.perf_counter()
for _ in range(K):
    acts.copy_(a_h, non_blocking=True); o_h.copy_(obs, non_blocking=True); r_h.copy_(rew, non_blocking=True)
    d_h.copy_(done, non_blocking=True); g_h.copy_(goal, non_blocking=True); torch.cuda.synchronize()
print(f"copies + sync alone: {1e3*(time.perf_counter()-t0)/K:.3f} ms/step")
