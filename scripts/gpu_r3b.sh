#!/bin/bash
mkdir -p gpurun_out
for ch in 256 512 1024 2048; do
  timeout 300 python bench.py --chunk $ch --steps 5 --warmup 3 --no-cpu --no-hamming 2>/dev/null | python -c "
import sys,json
for ln in sys.stdin:
    if ln.startswith('{'):
        d=json.loads(ln); print('chunk $ch', round(d['value']), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value']), {k:round(v['ms_per_step'],2) for k,v in d['stages'].items()})
"
done
