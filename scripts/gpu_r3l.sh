#!/bin/bash
timeout 600 python -m pytest tests/test_match_gpu.py tests/test_match2_gpu.py -m gpu -x -q 2>&1 | tail -3
timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu --no-hamming 2>/dev/null | python -c "
import sys,json
for ln in sys.stdin:
    if ln.startswith('{'):
        d=json.loads(ln); print(round(d['value']), json.dumps(d['e2e']))
"
