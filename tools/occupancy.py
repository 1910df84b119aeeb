"""Blocks per SM of the fp32 phase kernels from the library's own numbers (no GPU needed): dynamic shared memory per block
(adroit_debug_layout) + 1 KB reserved per block against 228 KB per SM, allocation granularity 128 B; registers from cuobjdump.
Usage: python tools/occupancy.py [lib.so]"""
import ctypes, os, re, subprocess, sys
import numpy as np
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(R, "mj_envs_b200", "libadroit_b200.so")
L = ctypes.CDLL(lib)
v = np.zeros(64, dtype=np.int32)
n = L.adroit_debug_layout(v.ctypes.data_as(ctypes.c_void_p), 64)
names = ["smooth", "collide", "solve<16>", "solve<big>", "finish", "reset"]
smem = dict(zip(names, map(int, v[27:33])))
regs = {}
out = subprocess.run(["cuobjdump", "-res-usage", lib], capture_output=True, text=True).stdout
for fn, rg in re.findall(r"Function (\S+):\s*\n?\s*REG:(\d+)", out):
    for k, pat in (("smooth", "smooth_kernelIf"), ("collide", "collide_kernelIf"), ("solve<16>", "solve_kernelIfLi16"), ("solve<big>", "solve_kernelIfLi100"),
                   ("finish", "finish_kernelIf"), ("reset", "reset_kernelIf")):
        if pat in fn: regs[k] = int(rg)
print(f"NVP {v[0]}  NTRI {v[6]}  context block {int(v[33]) * 4 / 1024:.1f} KB + {int(v[34]) * 4 / 1024:.1f} KB ints per env (fp32)")
for k in names:
    s = smem[k]
    a = (s + 127) // 128 * 128 + 1024
    bs = (228 * 1024) // a
    r = regs.get(k, 0)
    br = 65536 // (((r + 7) // 8 * 8) * 128) if r else 0
    print(f"{k:10s} smem/block {s:6d} B -> {bs:2d} blocks/SM   regs {r:3d} -> {br:2d} blocks/SM (at 4 warps per block; the small solve class runs 5)   => {min(bs, br) if r else bs} blocks")
