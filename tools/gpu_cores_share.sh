# does leaving room beside the solve kernel (3 resident blocks per SM instead of 4) let the other sub-batch streams' kernels fill its idle issue slots?
export WORKLOADS="hammer-v0:8192" STEPS=60 E2E=1
for cfg in "0 3" "2100 3" "2100 4" "2100 6" "0 4" "20000 3" "20000 6"; do set -- $cfg; echo "== solve extra smem $1, sub-batches $2"; env ADROIT_EXTRA_SMEM_2=$1 ADROIT_SUBBATCHES=$2 PHASE_TAG=share bash tools/gpu_phases.sh; done
