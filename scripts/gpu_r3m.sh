#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_extract_gpu.py tests/test_frame_gpu.py tests/test_host_api.py -m gpu -x -q 2>&1 | tail -3
timeout 600 python scripts/soak_parity.py --trials 1200 --seed 21 --budget-s 100 2>&1 | tail -2
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --no-hamming 2>/dev/null | python -c "
import sys,json
for ln in sys.stdin:
    if ln.startswith('{'):
        d=json.loads(ln); print(round(d['value']), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value']), {k:round(v['ms_per_step'],2) for k,v in d['stages'].items()})
"
