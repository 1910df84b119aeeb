"""Static SASS instruction count per source function for one kernel of an nvdisasm --print-line-info-inline dump.
Usage: python tools/code_size.py dis.txt kernel_substring [top]"""
import re, collections, os, sys
src_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "mj_envs_b200", "csrc")
fmap = {}
def funcs(f):
    if f in fmap: return fmap[f]
    p = os.path.join(src_dir, f); m = {}; name = "?"
    if os.path.exists(p):
        for i, t in enumerate(open(p), 1):
            mm = re.match(r"^(?:template <[^>]*>\s*)?(?:__device__|DI|PHASE_FN|SUB_FN|__global__|static|inline)[^;{]*?\b([A-Za-z_0-9]+)\s*\(", t)
            if mm and not t.startswith(" "): name = mm.group(1)
            m[i] = name
    fmap[f] = m; return m
want = sys.argv[2]; top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
inside = False; loc = None; fresh = True; cnt = collections.Counter()
for l in open(sys.argv[1]):
    if l.startswith(".text."): inside = want in l
    if not inside: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        if fresh:   # the first line of a group is the innermost (inlined) location
            f = os.path.basename(m.group(1)); loc = (f, funcs(f).get(int(m.group(2)), "?")); fresh = False
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l) and loc: cnt[loc] += 1; fresh = True
tot = sum(cnt.values()); print("total instructions", tot, "=", tot * 16 // 1024, "KB")
for k, v in cnt.most_common(top): print(f"{v:6d} {100*v/tot:5.1f}% {k[0]} {k[1]}")
