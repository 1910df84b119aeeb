# sub-batch pipelining A/B: parity tests with the default, then bench door/hammer for 1,2,4,8 sub-batches
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
for nsub in 1 2 4 8; do
  echo "== ADROIT_SUBBATCHES=$nsub"
  ADROIT_SUBBATCHES=$nsub timeout 600 python bench.py --steps 60 --warmup 10 --no-cpu-baseline --also hammer-v0:8192,pen-v0:16384 2>&1 | tail -1 > gpurun_out/sub_$nsub.json
  python -c "
import json,sys
d=json.loads(open('gpurun_out/sub_$nsub.json').read()); print('door', round(d['value']), 'e2e', round(d['e2e']['value']), 'also', {k: round(v['value']) for k,v in d.get('also',{}).items()}, 'launches', d['gpu_launches'])"
done
