for k in solve finish smooth collide; do
timeout 900 ncu --set full --clock-control none --import-source on -k regex:${k}_kernel -s 250 -c 1 -o gpurun_out/r2_${k}_hammer python bench.py --workload hammer-v0:8192 --steps 60 --warmup 3 --no-cpu-baseline --also '' > gpurun_out/r2_ncu_${k}.log 2>&1; echo rc=$?
done
