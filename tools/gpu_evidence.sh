# One GPU evidence run, parameterised:   gpurun --timeout 3000 -- 'bash tools/gpu_evidence.sh TAG [parts]'
# parts (default "tests smoke bench ref launches ncu"): tests = pytest -m gpu; smoke; bench = default line (all `also` legs + CPU arm);
# ref = --impl reference; launches = ncu launch list of the default workload; ncu = one `ncu --set full` capture per phase kernel
# (hammer-v0 x8192, steady state, ADROIT_SUBBATCHES=1 so that the captured launch is the whole 8192-env batch: a 512-block sub-batch
# launch never fills the machine and hides what the kernel is bound by at full occupancy) + source pages.  Artefacts land in gpurun_out/TAG_*; tools/collect_profiles.py TAG copies
# the judged ones into profiles/.
TAG=${1:-run}; PARTS=${2:-"tests smoke bench ref launches ncu"}
has() { case " $PARTS " in *" $1 "*) return 0;; *) return 1;; esac; }
set -x
if has tests; then timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest_gpu.log; tail -5 gpurun_out/${TAG}_pytest_gpu.log; fi
if has smoke; then timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/${TAG}_smoke.log; fi
if has bench; then timeout 1500 python bench.py > gpurun_out/${TAG}_bench_hammer.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"; cut -c1-600 gpurun_out/${TAG}_bench_hammer.json; fi
if has bench20; then timeout 900 python bench.py --steps 20 --warmup 5 --also none --no-cpu-baseline > gpurun_out/${TAG}_bench_hammer_20steps.json 2>> gpurun_out/${TAG}_bench.err; echo "bench20 rc=$?"; cut -c1-300 gpurun_out/${TAG}_bench_hammer_20steps.json; fi
if has ref; then timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_reference.json 2>/dev/null; echo "ref rc=$?"; cut -c1-400 gpurun_out/${TAG}_bench_reference.json; fi
B="python bench.py --steps 30 --warmup 3 --no-cpu-baseline --also none --e2e-steps 1"
if has launches; then
M=gpu__time_duration.sum,launch__registers_per_thread,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,launch__occupancy_limit_shared_mem,launch__occupancy_limit_registers,dram__bytes_read.sum,dram__bytes_write.sum
timeout 900 ncu --metrics $M --clock-control none -s 25000 -c 400 --csv --log-file gpurun_out/${TAG}_hammer_kernels.csv $B > gpurun_out/${TAG}_hammer_kernels.log 2>&1; echo rc=$?
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 25000 -c 600 --csv --log-file gpurun_out/${TAG}_launches.csv $B > gpurun_out/${TAG}_launches.log 2>&1; echo rc=$?
fi
if has ncu; then
for k in ${NCU_KERNELS:-solve finish smooth collide}; do
ADROIT_SUBBATCHES=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:${k}_kernel -s ${NCU_SKIP:-1200} -c 1 -o gpurun_out/${TAG}_${k}_hammer $B > gpurun_out/${TAG}_ncu_${k}.log 2>&1; echo rc=$?
timeout 300 ncu -i gpurun_out/${TAG}_${k}_hammer.ncu-rep --page source --csv --print-source cuda,sass > gpurun_out/${TAG}_${k}_hammer.source.csv 2>/dev/null; echo src rc=$?
done
fi
