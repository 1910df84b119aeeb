"""Executed SASS opcode mix, and the source lines behind the expensive sequences (MUFU.*, slow-path CALLs), from an
`ncu --page source --csv --print-source cuda,sass` export.  Usage: python tools/ncu_ops.py export.csv [top]"""
import csv, collections, os, re, sys
path = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 20
cur = None; line = (None, 0, ""); ops = collections.Counter(); hot = collections.Counter(); tot = 0
for r in csv.reader(open(path)):
    if not r: continue
    if r[0] == "File Path": cur = os.path.basename(r[1]); continue
    if r[0] in ("Line No", "Function Name") or len(r) < 8: continue
    if r[0].isdigit(): line = (cur, int(r[0]), r[1].strip()[:90]); continue      # a CUDA source line; its SASS rows follow
    if r[0] != "" or not r[2].startswith("0x"): continue
    try: ex = int(float(r[7] or 0))
    except ValueError: continue
    m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_]+(\.[A-Z0-9_]+)?)", r[3])
    if not m: continue
    op = m.group(2) if m.group(2).startswith("MUFU") else m.group(2).split(".")[0]
    ops[op] += ex; tot += ex
    if op.startswith("MUFU") or op in ("CALL", "RET"): hot[(line[0], line[1], op, line[2])] += ex
print("total warp instructions", tot)
for op, n in ops.most_common(top): print(f"  {op:12s} {n:12d} {100 * n / max(tot, 1):5.1f}%")
print("-- expensive sequences by source line")
for (f, ln, op, src), n in hot.most_common(top): print(f"  {n:10d} {op:10s} {f}:{ln}  {src}")
