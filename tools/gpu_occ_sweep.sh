# phase-kernel throughput against resident blocks per SM (unused dynamic shared memory forces fewer resident blocks), hammer-v0 x8192
# usage: bash tools/gpu_occ_sweep.sh PHASE "BYTES..."    (phase 0 smooth, 1 collide, 2 solve, 3 finish)
export WORKLOADS="hammer-v0:8192" STEPS=40 E2E=1
P=${1:-2}
for x in ${2:-0 12000 30000 60000}; do echo "== phase $P extra smem $x"; env ADROIT_EXTRA_SMEM_$P=$x PHASE_TAG=p${P}x$x bash tools/gpu_phases.sh; done
