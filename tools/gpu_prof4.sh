# ncu --set full of one late launch of each phase kernel (hammer-v0 x 8192), source-level export for tools/ncu_lines.py
for k in ${KERNELS:-solve finish}; do
timeout 900 ncu --set full --clock-control none --import-source on -k regex:${k}_kernel -s 250 -c 1 -o gpurun_out/r3_${k}_hammer python bench.py --workload hammer-v0:8192 --steps 60 --warmup 3 --no-cpu-baseline --also '' > gpurun_out/r3_ncu_${k}.log 2>&1; echo rc=$?
done
