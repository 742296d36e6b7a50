#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_match_gpu.py tests/test_bow_gpu.py tests/test_kfmatch_gpu.py tests/test_host_api.py -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r2v_pytest.log
ORBX_LAT_TRACE=1 timeout 600 python bench.py --workload matcher_calls --steps 2 --no-cpu 2>&1 >/dev/null | grep "orbx lat" | sort -u -k3,3
timeout 600 python bench.py --workload matcher_calls --steps 2 > gpurun_out/r2v_matcher_calls.json 2> gpurun_out/r2v_matcher_calls.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2v_matcher_calls.json'))
for k,v in d['calls'].items(): print(k, v['gpu_us'], v.get('cpu_oracle_us'), v.get('speedup'))
PY
