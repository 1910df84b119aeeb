export STEPS=100 E2E=100
for w in "door-v0:4096 1" "door-v0:4096 2" "door-v0:4096 3" "pen-v0:16384 3" "pen-v0:16384 4" "pen-v0:16384 6" "relocate-v0:8192 3" "relocate-v0:8192 4"; do set -- $w; echo "== $1 subbatches $2"; WORKLOADS=$1 ADROIT_SUBBATCHES=$2 PHASE_TAG=sb$2 bash tools/gpu_phases.sh | cut -c1-120; done
