set -x
nvidia-smi -L
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r1_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1_pytest_gpu.log
tail -5 gpurun_out/r1_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r1_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r1_smoke.log
timeout 600 python bench.py > gpurun_out/r1_bench.json 2> gpurun_out/r1_bench.err; echo "bench rc=$?"; cat gpurun_out/r1_bench.json
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r1_bench_ref.json 2> gpurun_out/r1_bench_ref.err; echo "ref rc=$?"; cat gpurun_out/r1_bench_ref.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1_launches.csv python bench.py --steps 10 --warmup 3 --no-cpu-baseline --also '' > gpurun_out/r1_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:env_kernel -s 8 -c 2 -o gpurun_out/r1_env_kernel_door python bench.py --steps 10 --warmup 3 --no-cpu-baseline --also '' > gpurun_out/r1_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out
