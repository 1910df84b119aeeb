"""Opcode histogram of the shipped cubin per kernel (no GPU): cuobjdump -xelf + nvdisasm.  Usage: python tools/sass_opcodes.py [lib.so] [--all]
Default: the fp32 phase kernels; writes nothing, prints a table (profiles/sass_opcodes.txt is this output)."""
import collections, os, re, subprocess, sys, tempfile
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
args = [a for a in sys.argv[1:] if not a.startswith("--")]
lib = args[0] if args else os.path.join(R, "mj_envs_b200", "libadroit_b200.so")
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, capture_output=True)
cub = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")][0]
txt = subprocess.run(["nvdisasm", "-c", cub], capture_output=True, text=True).stdout
sec = None; H = collections.defaultdict(collections.Counter)
for line in txt.splitlines():
    m = re.match(r"\s*\.section\s+\.text\.(\S+?),", line)
    if m: sec = m.group(1); continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_]*)((?:\.[A-Z0-9_]+)*)", line)
    if sec and m: H[sec][m.group(1) + ("" if m.group(1) not in ("UBLKCP", "LDGSTS", "SYNCS", "LD", "ST", "ATOMS", "RED") else m.group(2))] += 1
want = ["smooth_kernelIf", "collide_kernelIf", "solve_kernelIfLi16", "solve_kernelIfLi100", "finish_kernelIf", "reset_kernelIf"]
for s in sorted(H):
    if "--all" not in sys.argv and not any(w in s for w in want): continue
    tot = sum(H[s].values())
    name = re.sub(r"^_ZN6adroit\d+", "", s)[:34]
    key = {k: H[s].get(k, 0) for k in ("LDS", "STS", "LDG", "STG", "LDL", "STL", "SHFL", "CALL", "BAR", "FFMA", "MUFU")}
    gen = sum(v for k, v in H[s].items() if k.startswith("LD.") or k == "LD" or k.startswith("ST.") or k == "ST")
    bulk = {k: v for k, v in H[s].items() if k.startswith(("UBLKCP", "LDGSTS", "SYNCS", "UTMA"))}
    print(f"{name:34s} {tot:6d} instr = {tot * 16 / 1024:6.1f} KB | " + " ".join(f"{k} {v}" for k, v in key.items()) + f" | generic LD/ST {gen} | {bulk}")
