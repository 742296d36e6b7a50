#!/bin/bash
mkdir -p gpurun_out
for wl in hd stereo rgbd knn; do
  timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu > gpurun_out/r02z_bench_$wl.json 2> gpurun_out/r02z_bench_$wl.err; echo "$wl rc=$?"
  python - $wl <<'PY'
import json,sys
try:
    d=json.loads(open(f'gpurun_out/r02z_bench_{sys.argv[1]}.json').read().strip().splitlines()[-1])
    print(sys.argv[1], d['metric'][:60], 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'ms/step %.2f'%d['ms_per_step'])
except Exception as e: print('parse fail', e); print(open(f'gpurun_out/r02z_bench_{sys.argv[1]}.err').read()[-600:])
PY
done
