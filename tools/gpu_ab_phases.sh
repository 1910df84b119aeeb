# per-phase A/B: every library variant (product + build_ab/*.so) on the workloads in $WORKLOADS
for lib in mj_envs_b200/libadroit_b200.so build_ab/*.so; do
  [ -f "$lib" ] || continue
  echo "== $lib"
  ADROIT_B200_LIB=$PWD/$lib PHASE_TAG=$(basename $lib .so) bash tools/gpu_phases.sh
done
