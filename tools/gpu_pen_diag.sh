WORKLOADS="door-v0:16384 pen-v0:4096" bash tools/gpu_ab_phases.sh
for v in new old; do
  lib=mj_envs_b200/libadroit_b200.so; [ $v = old ] && lib=build_ab/old.so
  ADROIT_B200_LIB=$PWD/$lib timeout 600 ncu --set full --clock-control none --import-source on -k regex:solve_kernel -s 100 -c 1 -o gpurun_out/pen_solve_$v python bench.py --workload pen-v0:16384 --steps 30 --warmup 3 --no-cpu-baseline --also '' > gpurun_out/pen_ncu_$v.log 2>&1; echo rc=$?
done
