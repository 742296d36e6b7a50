mkdir -p gpurun_out
for c in 64 128 256 512 1024; do
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --chunk $c > gpurun_out/chunk_$c.json 2> gpurun_out/chunk_$c.err
python - $c <<'PY'
import json,sys
d=json.load(open(f'gpurun_out/chunk_{sys.argv[1]}.json'))
print('chunk',sys.argv[1],'value',round(d['value']),'e2e',round(d['e2e']['value']),'ms/step',round(d['ms_per_step'],2), {k:round(v['ms_per_step'],2) for k,v in d['stages'].items()})
PY
done
