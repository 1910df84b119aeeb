"""Static code size and instruction-fetch stalls per source function from an `ncu --page source --csv --print-source cuda,sass` export.
Usage: python tools/ncu_codesize.py export.csv"""
import csv, collections, os, re, sys
path = sys.argv[1]
cur = None; hdr = None; curline = None
rows = []   # (file, line, addr, sass, samples, executed, no_inst)
for r in csv.reader(open(path)):
    if not r: continue
    if r[0] == "File Path": cur = os.path.basename(r[1]); continue
    if r[0] == "Line No": hdr = r; continue
    if r[0] == "Function Name" or hdr is None: continue
    d = dict(zip(hdr[4:], r[4:]))
    def g(k):
        try: return int(float(d.get(k) or 0))
        except Exception: return 0
    if r[0].isdigit(): curline = int(r[0]); continue
    if r[0] == "" and r[2].startswith("0x"):
        rows.append((cur, curline, int(r[2], 16), r[3].strip(), g("# Samples"), g("Instructions Executed"), g("stall_no_inst")))
src_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "mj_envs_b200", "csrc")
fmap = {}
for f in set(x[0] for x in rows):
    p = os.path.join(src_dir, f)
    if not os.path.exists(p): continue
    name = "?"; m = {}
    for i, t in enumerate(open(p), 1):
        mm = re.match(r"^(?:template <[^>]*>\s*)?(?:__device__|DI|PHASE_FN|SUB_FN|__global__|static|inline)[^;{]*?\b([A-Za-z_0-9]+)\s*\(", t)
        if mm and not t.startswith(" "): name = mm.group(1)
        m[i] = name
    fmap[f] = m
F = collections.defaultdict(lambda: [0, 0, 0, 0])
for f, ln, addr, sass, smp, ex, ni in rows:
    k = (f, fmap.get(f, {}).get(ln, "?"))
    F[k][0] += 1; F[k][1] += smp; F[k][2] += ex; F[k][3] += ni
n = len(rows); tot_s = sum(x[4] for x in rows); tot_e = sum(x[5] for x in rows); tot_n = sum(x[6] for x in rows)
print(f"SASS instructions {n} = {n * 16 / 1024:.1f} KB; samples {tot_s}; executed {tot_e}; no_inst samples {tot_n} ({100 * tot_n / max(tot_s, 1):.1f} %)")
# hot working set: smallest set of 128-byte lines covering 90 / 99 % of executed instructions
lines = collections.Counter()
for f, ln, addr, sass, smp, ex, ni in rows: lines[addr >> 7] += ex
acc = 0; cnt = 0; marks = {}
for l, e in lines.most_common():
    acc += e; cnt += 1
    for q in (0.5, 0.9, 0.99):
        if q not in marks and acc >= q * tot_e: marks[q] = cnt
print("128-byte code lines covering 50/90/99 % of executed instructions:", {q: f"{c} lines = {c / 8:.1f} KB" for q, c in marks.items()})
print(f"{'file':18s} {'function':24s} {'code KB':>8s} {'exec %':>7s} {'samples %':>9s} {'no_inst %':>9s}")
for k, v in sorted(F.items(), key=lambda kv: -kv[1][0])[:45]:
    print(f"{k[0]:18s} {k[1]:24s} {v[0] * 16 / 1024:8.2f} {100 * v[2] / max(tot_e, 1):7.2f} {100 * v[1] / max(tot_s, 1):9.2f} {100 * v[3] / max(tot_n, 1):9.2f}")
