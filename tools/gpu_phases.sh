# per-phase device times of the four tasks (bench.py roofline.phase_ms_per_launch)
for w in ${WORKLOADS:-door-v0:4096 hammer-v0:8192 pen-v0:16384 relocate-v0:8192}; do
  timeout 600 python bench.py --workload $w --steps ${STEPS:-100} --warmup 10 --no-cpu-baseline --also '' 2>&1 | tail -1 > gpurun_out/phase_${w%%:*}.json
  python -c "
import json
d=json.load(open('gpurun_out/phase_${w%%:*}.json')); r=d['roofline']
print('$w', round(d['value']), 'e2e', round(d['e2e']['value']), {k: round(v,4) for k,v in r['phase_ms_per_launch'].items()}, 'ms/step', round(d['ms_per_step'],3))"
done
