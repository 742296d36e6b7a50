# chunk-size sweep of the default workload (run under gpurun)
mkdir -p gpurun_out
for c in ${@:-128 256 384 512 1024}; do
  timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --chunk $c > gpurun_out/chunk_$c.json 2> gpurun_out/chunk_$c.err
  python - $c <<'PY'
import json,sys
d=json.loads(open(f'gpurun_out/chunk_{sys.argv[1]}.json').read().strip().splitlines()[-1])
print('chunk',sys.argv[1],'value',round(d['value']),'e2e',round(d['e2e']['value']),{k:round(v['ms_per_step'],2) for k,v in d['stages'].items()})
PY
done
