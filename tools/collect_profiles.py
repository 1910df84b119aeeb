"""Copies the judged artefacts of a GPU evidence run from gpurun_out/ into profiles/ and writes profiles/README.md.
Usage: python tools/collect_profiles.py r2"""
import glob, json, os, shutil, subprocess, sys
tag = sys.argv[1]
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src, dst = os.path.join(R, "gpurun_out"), os.path.join(R, "profiles")
lines = [f"# profiles/ — evidence of run `{tag}` (see DESIGN.md §5)\n",
         "Evidence runs are made by `tools/gpu_evidence.sh TAG` (one `gpurun` call each); tags r1…r9 = build round 1, R2* = build round 2. "
         f"`{tag}` is the current state of the code, the older files are kept as the history DESIGN.md §5 quotes. Each run: "
         "`pytest -m gpu`, `smoke()`, `bench.py` on the default workload (with the CPU arm) and on the other three tasks + the pixel "
         "config, the ncu launch lists (`--metrics gpu__time_duration.sum … --clock-control none`), and one `ncu --set full` capture "
         "per phase kernel (summaries by `tools/ncu_summary.py`; per-line / per-opcode views by `tools/ncu_lines.py`, `tools/ncu_ops.py`).\n"]
traffic = {}
for f in sorted(glob.glob(os.path.join(src, f"{tag}_bench_*.json"))):
    shutil.copy(f, dst)
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception:
        continue
    rf = d.get("roofline") or {}
    lines.append(f"* `{os.path.basename(f)}`: {d.get('config', {}).get('workload', '')[:60]} — **{d['value']:.3e} {d['unit']}**, "
                 f"e2e {d.get('e2e', {}).get('value', float('nan')):.3e}, {d['ms_per_step']:.3f} ms/step, launches {d.get('gpu_launches')}, "
                 f"dominant kernel {rf.get('kernel')} ({100 * (rf.get('kernel_share_of_step') or 0):.0f} % of the step), "
                 f"roofline frac {rf.get('frac', float('nan')):.2e}" + (f", cpu_baseline {d['cpu_baseline']['value']:.3e} on {d['cpu_baseline']['cores']} cores" if d.get("cpu_baseline") else ""))
for f in glob.glob(os.path.join(src, f"{tag}_*.ncu-rep")):
    name = os.path.basename(f)[:-8]
    out = subprocess.run([sys.executable, os.path.join(R, "tools", "ncu_summary.py"), f], capture_output=True, text=True).stdout
    open(os.path.join(dst, name + ".summary.txt"), "w").write(out)
    rd = wr = None
    for l in out.splitlines():
        p = l.split()
        if "dram__bytes_read.sum" in l: rd = float(p[1]) * (1e6 if p[2].startswith("M") else 1e9 if p[2].startswith("G") else 1e3 if p[2].startswith("K") else 1)
        if "dram__bytes_write.sum" in l: wr = float(p[1]) * (1e6 if p[2].startswith("M") else 1e9 if p[2].startswith("G") else 1e3 if p[2].startswith("K") else 1)
    grid = None; wpb = 4
    for l in out.splitlines():
        if "launch__grid_size" in l: grid = float(l.split()[1])
        if "launch__block_size" in l: wpb = int(float(l.split()[1])) // 32
    task = name.rsplit("_", 1)[-1]
    if rd is not None and grid and "solve" in name:
        # DRAM bytes per env of ONE launch of the dominant kernel (block size / 32 envs per block); bench.py scales it to its launch size
        traffic[f"{task}-v0"] = {"dram_bytes_per_env_launch": (rd + wr) / (grid * wpb), "kernel": "solve_kernel", "envs_in_captured_launch": int(grid * wpb),
                                 "source": f"profiles/{name}.summary.txt"}
    lines.append(f"* `{name}.summary.txt`: `ncu --set full` of one launch ({task}-v0, late episode)")
for f, tool in ((f"{tag}_hammer_kernels.csv", "ncu_kernels.py"), (f"{tag}_launches.csv", "ncu_launches.py")):
    p = os.path.join(src, f)
    if os.path.exists(p):
        shutil.copy(p, dst)
        out = subprocess.run([sys.executable, os.path.join(R, "tools", tool), p], capture_output=True, text=True).stdout
        open(os.path.join(dst, f.replace(".csv", ".txt")), "w").write(out)
        lines.append(f"* `{f}` / `.txt`: launch list ({tool})")
for f in (f"{tag}_pytest_gpu.log", f"{tag}_smoke.log"):
    if os.path.exists(os.path.join(src, f)): shutil.copy(os.path.join(src, f), dst); lines.append(f"* `{f}`")
tp = os.path.join(dst, "traffic.json")
old = json.load(open(tp)) if os.path.exists(tp) else {}
if traffic:
    old.update(traffic)
    old = {k: v for k, v in old.items() if isinstance(v, dict)}
    old.update(traffic)
    old["_source"] = "dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the dominant kernel (solve_kernel), ncu --set full, per env of that launch"
    json.dump(old, open(tp, "w"))
open(os.path.join(dst, "README.md"), "w").write("\n".join(lines) + "\n")
print("\n".join(lines))
