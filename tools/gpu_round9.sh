# Round-2 evidence run: tests, smoke, benches (both arms), per-kernel launch lists, one full ncu capture per phase kernel.
set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r9_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r9_pytest_gpu.log; tail -3 gpurun_out/r9_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r9_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r9_smoke.log
timeout 900 python bench.py > gpurun_out/r9_bench_door4096.json 2> gpurun_out/r9_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r9_bench_reference.json 2>/dev/null; echo "ref rc=$?"
for w in hammer-v0:8192 pen-v0:16384 relocate-v0:8192; do t=${w%%-*}; timeout 600 python bench.py --workload $w --steps 100 --warmup 20 --no-cpu-baseline --also '' > gpurun_out/r9_bench_$t.json 2>/dev/null; echo "$t rc=$?"; done
M=gpu__time_duration.sum,launch__registers_per_thread,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,launch__occupancy_limit_shared_mem,launch__occupancy_limit_registers,dram__bytes_read.sum,dram__bytes_write.sum
timeout 900 ncu --metrics $M --clock-control none -s 1300 -c 200 --csv --log-file gpurun_out/r9_hammer_kernels.csv python bench.py --workload hammer-v0:8192 --steps 70 --warmup 3 --no-cpu-baseline --also '' > gpurun_out/r9_hammer_kernels.log 2>&1; echo rc=$?
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r9_launches.csv python bench.py --steps 10 --warmup 3 --no-cpu-baseline --also '' > gpurun_out/r9_launches.log 2>&1; echo rc=$?
for k in solve finish smooth collide; do
timeout 900 ncu --set full --clock-control none --import-source on -k regex:${k}_kernel -s 250 -c 1 -o gpurun_out/r9_${k}_hammer python bench.py --workload hammer-v0:8192 --steps 60 --warmup 3 --no-cpu-baseline --also '' > gpurun_out/r9_ncu_${k}.log 2>&1; echo rc=$?
done
timeout 600 python bench.py --workload hammer-v0:8192 --pixels --steps 40 --warmup 5 --no-cpu-baseline --also '' > gpurun_out/r9_bench_hammer_pixels.json 2>/dev/null; echo "pixels rc=$?"
for k in solve finish smooth; do
timeout 300 ncu -i gpurun_out/r9_${k}_hammer.ncu-rep --page source --csv --print-source cuda,sass > gpurun_out/r9_${k}_hammer.source.csv 2>/dev/null; echo src rc=$?
done
# the default workload's dominant kernel (door-v0 x 4096): one full capture for roofline.traffic
timeout 900 ncu --set full --clock-control none --import-source on -k regex:solve_kernel -s 60 -c 1 -o gpurun_out/r9_solve_door python bench.py --steps 60 --warmup 3 --no-cpu-baseline --also '' > gpurun_out/r9_ncu_solve_door.log 2>&1; echo rc=$?
