set -x
for t in door hammer pen relocate; do timeout 300 python tools/gpu_stats.py --task $t --n 2048 --steps 120 > gpurun_out/r1b_stats_$t.log 2>&1; tail -5 gpurun_out/r1b_stats_$t.log; done
for w in hammer-v0:8192 pen-v0:16384 relocate-v0:8192 door-v0:16384; do timeout 600 python bench.py --workload $w --steps 100 --warmup 20 --no-cpu-baseline --also '' > gpurun_out/r1b_bench_${w%%:*}.json 2>&1; cat gpurun_out/r1b_bench_${w%%:*}.json; done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:env_kernel -s 150 -c 1 -o gpurun_out/r1b_env_kernel_door_late python bench.py --steps 200 --warmup 3 --no-cpu-baseline --also '' > gpurun_out/r1b_ncu_door.log 2>&1; echo "rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:env_kernel -s 150 -c 1 -o gpurun_out/r1b_env_kernel_hammer_late python bench.py --workload hammer-v0:8192 --steps 200 --warmup 3 --no-cpu-baseline --also '' > gpurun_out/r1b_ncu_hammer.log 2>&1; echo "rc=$?"
