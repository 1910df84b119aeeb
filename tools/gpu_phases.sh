# per-phase device times (bench.py roofline.phase_ms_per_launch) of the workloads in $WORKLOADS, steady state
for w in ${WORKLOADS:-hammer-v0:8192 door-v0:4096 pen-v0:16384 relocate-v0:8192}; do
  f=gpurun_out/phase_${w%%:*}${PHASE_TAG:+_$PHASE_TAG}.json
  timeout 600 python bench.py --workload $w --steps ${STEPS:-60} --warmup 5 --no-cpu-baseline --also none --e2e-steps ${E2E:-60} 2>&1 | tail -1 > $f
  python -c "
import json
d=json.load(open('$f')); r=d['roofline']
print('$w', round(d['value']), 'e2e', round(d['e2e']['value']), 'api', round(d['e2e_api']['value']), {k[:-7]: round(v,4) for k,v in r['phase_ms_per_launch'].items()}, 'ms/step', round(d['ms_per_step'],3), 'launches/step', d['gpu_launches']/d['steps'])"
done
