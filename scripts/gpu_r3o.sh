#!/bin/bash
timeout 900 python -m pytest tests/test_extract_gpu.py tests/test_match_gpu.py tests/test_host_api.py -m gpu -x -q 2>&1 | tail -3
timeout 600 python scripts/soak_parity.py --trials 1200 --seed 31 --budget-s 90 2>&1 | tail -2
for t in 8 4 16; do
ORBX_HOST_THREADS=$t timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu --no-hamming 2>/dev/null | python -c "
import sys,json
for ln in sys.stdin:
    if ln.startswith('{'):
        d=json.loads(ln); print('threads $t', round(d['value']), 'e2e', round(d['e2e']['value']), 'pageable', round(d['e2e']['pageable_buffers_frames_s_rank0']))
"
done
