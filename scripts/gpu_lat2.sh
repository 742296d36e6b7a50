# usage: gpu_lat2.sh VAR=val ... (run under gpurun) — graph parity test + latency line per env variant
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_extract_gpu.py -m gpu -x -q 2>&1 | tail -2
for v in "$@"; do echo "== $v"; env $v timeout 300 python bench.py --workload single --no-cpu --steps 2 2>&1 | python -c "
import sys,json
for ln in sys.stdin:
    if ln.startswith('{'):
        d=json.loads(ln)['latency_us']; print({k: round(x,1) for k,x in d.items()})
    else: print(ln.rstrip()[-300:])
"; done
