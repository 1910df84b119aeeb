"""Static SASS size of one kernel by source function, from the built library (no GPU): cuobjdump -xelf + nvdisasm -g.
Usage: python tools/sass_size.py [lib.so] [kernel-substring, default solve_kernelIfLi16]   (device functions the kernel calls
(`SUB_FN`) live in their own .text sections and are listed separately with --all)"""
import collections, os, re, subprocess, sys, tempfile
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
args = [a for a in sys.argv[1:] if not a.startswith("--")]
lib = args[0] if args and args[0].endswith(".so") else os.path.join(R, "mj_envs_b200", "libadroit_b200.so")
pat = [a for a in args if not a.endswith(".so")]
pat = pat[0] if pat else "solve_kernelIfLi16"
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, capture_output=True)
cub = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")][0]
txt = subprocess.run(["nvdisasm", "-gi", "-c", cub], capture_output=True, text=True).stdout
src_dir = os.path.join(R, "mj_envs_b200", "csrc")
fmap = {}
def fn_of(f, ln):
    b = os.path.basename(f)
    if b not in fmap:
        p = os.path.join(src_dir, b); m = {}
        if os.path.exists(p):
            name = "?"
            for i, t in enumerate(open(p), 1):
                mm = re.match(r"^(?:template <[^>]*>\s*)?(?:__device__|DI|PHASE_FN|SUB_FN|__global__|static|inline)[^;{]*?\b([A-Za-z_0-9]+)\s*\(", t)
                if mm and not t.startswith(" "): name = mm.group(1)
                m[i] = name
        fmap[b] = m
    return fmap[b].get(ln, b)
sec = None; chain = []; fresh = True; sizes = collections.defaultdict(collections.Counter); outer = collections.defaultdict(collections.Counter)
for line in txt.splitlines():
    m = re.match(r"\s*\.section\s+\.text\.(\S+?),", line)
    if m: sec = m.group(1); continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', line)
    if m:
        if fresh: chain = []; fresh = False
        chain.append((m.group(1), int(m.group(2)))); continue
    if sec and re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+\S", line):
        fresh = True
        repo = [c for c in chain if "/csrc/" in c[0]]
        inner = fn_of(*repo[0]) if repo else "?"
        # outermost repo frame below the kernel body itself (the phase function the code was inlined into)
        below = [c for c in repo if not (c[0].endswith("adroit_b200.cu") and fn_of(*c) in ("__launch_bounds__", "solve_kernel", "finish_kernel", "smooth_kernel", "collide_kernel", "reset_kernel"))]
        sizes[sec][inner] += 1
        outer[sec][fn_of(*below[-1]) if below else "(kernel body)"] += 1
secs = [s for s in sizes if pat in s]
tot_all = 0
for s in secs:
    tot = sum(sizes[s].values()); tot_all += tot
    print(f"== {s[:70]}: {tot} instructions = {tot * 16 / 1024:.1f} KB")
    print("   -- by the phase function the code was inlined into")
    for k, v in outer[s].most_common(20):
        print(f"   {k:28s} {v * 16 / 1024:7.2f} KB")
    print("   -- by innermost function")
    for k, v in sizes[s].most_common(40 if "--all" not in sys.argv else 200):
        print(f"   {k:28s} {v * 16 / 1024:7.2f} KB")
