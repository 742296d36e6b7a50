# usage: gpu_r2c.sh <tag>  (1 GPU) — all GPU tests, default bench (with the measured config-5 leg), int-rate probe, pipe metrics of the matcher family
T=${1:-r02c}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${T}_pytest.log
tail -6 gpurun_out/${T}_pytest.log
timeout 900 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
tail -3 gpurun_out/${T}_bench.err
python - $T <<'PY'
import json,sys
d=json.loads(open(f'gpurun_out/{sys.argv[1]}_bench.json').read().strip().splitlines()[-1])
print('value',round(d['value']),'e2e',round(d['e2e']['value']),'ms/step',round(d['ms_per_step'],2))
for k,v in d['stages'].items(): print(f"  {k:12s} {v['ms_per_step']:8.3f} ms  {v['launches_per_step']:.0f} launches")
print(json.dumps(d['hamming'],indent=1)); print(d['roofline']); print(d.get('cpu_baseline')); print('compares/step', d['hamming_compares_per_step'])
PY
