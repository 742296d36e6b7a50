# usage: gpu_lat.sh  (run under gpurun) — single-frame latency experiments: graph parity test, then the latency line under env variants
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_extract_gpu.py -m gpu -x -q 2>&1 | tail -3
run() { echo "== $*"; env "$@" timeout 300 python bench.py --workload single --no-cpu --steps 2 2>&1 | python -c "
import sys,json
for ln in sys.stdin:
    if ln.startswith('{'):
        d=json.loads(ln); print(d['latency_us'])
    else: print(ln.rstrip()[-300:])
"; }
run A=1
run ORBX_LAT_NOFORK=1
run ORBX_NO_GRAPH=1
for v in "$@"; do run $v; done
