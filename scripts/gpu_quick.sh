# quick check under gpurun: extractor parity tests + short bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_extract_gpu.py tests/test_match_gpu.py -m gpu -x -q > gpurun_out/quick_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/quick_pytest.log
tail -5 gpurun_out/quick_pytest.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu "$@" > gpurun_out/quick_bench.json 2> gpurun_out/quick_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/quick_bench.json'))
print('value',round(d['value']),'e2e',round(d['e2e']['value']),'ms/step',round(d['ms_per_step'],2))
for k,v in d['stages'].items(): print(f"  {k:12s} {v['ms_per_step']:8.3f} ms  {v['launches_per_step']:.0f} launches")
print(d['roofline'])
PY
