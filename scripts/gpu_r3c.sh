#!/bin/bash
mkdir -p gpurun_out
run() { # chunk hostchunk
  ORBX_HOST_CHUNK=$2 timeout 300 python bench.py --chunk $1 --steps 5 --warmup 3 --no-cpu --no-hamming 2>/dev/null | python -c "
import sys,json
for ln in sys.stdin:
    if ln.startswith('{'):
        d=json.loads(ln); print('chunk $1 host $2', round(d['value']), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value']))
"; }
run 1024 128; run 1024 192; run 1024 256; run 1024 320; run 2048 256; run 4096 256
