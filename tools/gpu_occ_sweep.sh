# solve-kernel throughput against resident blocks per SM (unused dynamic shared memory forces 5 -> 4 -> 3 -> 2 blocks), hammer-v0 x8192
export WORKLOADS="hammer-v0:8192" STEPS=60 E2E=1
for x in 0 12000 30000 60000; do echo "== extra smem $x"; ADROIT_EXTRA_SMEM_SOLVE=$x PHASE_TAG=x$x bash tools/gpu_phases.sh; done
