# A/B: bench each library variant in build_ab/ (plus the product library) on door and hammer
for lib in mj_envs_b200/libadroit_b200.so build_ab/*.so; do
  echo "== $lib"
  ADROIT_B200_LIB=$PWD/$lib timeout 600 python bench.py --steps 60 --warmup 10 --no-cpu-baseline --also hammer-v0:8192 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('door', round(d['value']), 'e2e', round(d['e2e']['value']), 'hammer', {k: round(v['value']) for k,v in d.get('also',{}).items()})"
done
