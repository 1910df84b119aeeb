"""Aggregates an `ncu --page source --csv --print-source cuda,sass` export by CUDA source line and by function.
Usage: python tools/ncu_lines.py export.csv [top_n]"""
import csv, collections, re, sys, os
path = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur = None; hdr = None; lines = collections.defaultdict(lambda: [0, 0, 0, 0, ""])
for r in csv.reader(open(path)):
    if not r: continue
    if r[0] == "File Path": cur = os.path.basename(r[1]); continue
    if r[0] == "Line No": hdr = r; continue
    if r[0] == "Function Name" or hdr is None or not r[0].isdigit(): continue
    d = dict(zip(hdr[4:], r[4:]))
    g = lambda k: int(float(d.get(k) or 0)) if (d.get(k) or "0").replace(".", "").isdigit() else 0
    L = lines[(cur, int(r[0]))]
    L[0] += g("# Samples"); L[1] += g("Instructions Executed"); L[2] += g("stall_no_inst"); L[3] += g("stall_long_sb"); L[4] = r[1]
tot = sum(v[0] for v in lines.values()); toti = sum(v[1] for v in lines.values())
print("total samples", tot, "warp instructions", toti)
# function map from the sources
funcs = collections.defaultdict(lambda: [0, 0, 0, 0])
src_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "mj_envs_b200", "csrc")
fmap = {}
for f in set(k[0] for k in lines):
    p = os.path.join(src_dir, f)
    if not os.path.exists(p): continue
    name = "?"; m = {}
    for i, t in enumerate(open(p), 1):
        mm = re.match(r"^(?:template <[^>]*>\s*)?(?:__device__|DI|PHASE_FN|SUB_FN|__global__|static|inline)[^;{]*?\b([A-Za-z_0-9]+)\s*\(", t)
        if mm and not t.startswith(" "): name = mm.group(1)
        m[i] = name
    fmap[f] = m
for (f, ln), v in lines.items():
    fn = fmap.get(f, {}).get(ln, "?")
    F = funcs[(f, fn)]
    for i in range(4): F[i] += v[i]
print("\n-- by function: samples%  instr%  no_inst  long_sb")
for k, v in sorted(funcs.items(), key=lambda kv: -kv[1][0])[:topn]:
    print(f"{k[0]:18s} {k[1]:24s} {100*v[0]/tot:6.2f}% {100*v[1]/max(toti,1):6.2f}% {v[2]:8d} {v[3]:8d}")
print("\n-- by line")
for k, v in sorted(lines.items(), key=lambda kv: -kv[1][0])[:topn]:
    print(f"{k[0]:16s}:{k[1]:4d} {100*v[0]/tot:6.2f}% inst {100*v[1]/max(toti,1):5.2f}% noinst {v[2]:6d} lsb {v[3]:6d} | {v[4][:100]}")
