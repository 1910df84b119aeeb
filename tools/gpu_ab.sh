# A/B of library variants and runtime switches on the steady-state phase timings:  gpurun -- 'bash tools/gpu_ab.sh "hammer-v0:8192 door-v0:4096"'
# every line: workload, env-steps/s (device), e2e, per-phase ms per launch.  Variants: build_ab/*.so + the product library;
# extra runs of the product library with environment switches: AB_ENVS="ADROIT_PDL=1 ADROIT_LIGHT_RESET=0"
export WORKLOADS=${1:-"hammer-v0:8192 door-v0:4096"}
export STEPS=${STEPS:-100} E2E=${E2E:-100}
run() { echo "== $1"; shift; env "$@" bash tools/gpu_phases.sh; }
for lib in build_ab/*.so; do [ -f "$lib" ] && run "$lib" ADROIT_B200_LIB=$PWD/$lib PHASE_TAG=$(basename $lib .so); done
run product PHASE_TAG=product
for e in $AB_ENVS; do run "product $e" $e PHASE_TAG=${e%%=*}; done
