# usage: gpu_check_all.sh <tag> — whole GPU suite, short soaks (extractor + matchers), matcher call latencies
T=${1:-chk}
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/${T}_pytest.log
timeout 600 python scripts/soak_parity.py --trials 1500 --seed 11 --budget-s 150 2>&1 | tail -2 | tee gpurun_out/${T}_soak.log
timeout 600 python scripts/soak_matchers.py --trials 8000 --seed 11 --budget-s 150 2>&1 | tail -2 | tee gpurun_out/${T}_soak_matchers.log
timeout 600 python bench.py --workload matcher_calls --steps 2 > gpurun_out/${T}_matcher_calls.json 2> gpurun_out/${T}_matcher_calls.err
python - $T <<'PY'
import json,sys
d=json.load(open(f'gpurun_out/{sys.argv[1]}_matcher_calls.json'))
for k,v in d['calls'].items(): print(k, v['gpu_us'], v.get('cpu_oracle_us'), v.get('speedup'))
PY
