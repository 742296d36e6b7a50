mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_match2_gpu.py -m gpu -x -q -k "stereo" > gpurun_out/stereo_pytest.log 2>&1; tail -3 gpurun_out/stereo_pytest.log
timeout 900 python bench.py --workload stereo --steps 3 --warmup 3 > gpurun_out/bench_stereo.json 2> gpurun_out/bench_stereo.err; tail -2 gpurun_out/bench_stereo.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_stereo.json').read().strip().splitlines()[-1])
print('value %.4g'%d['value'],'e2e %.4g'%d['e2e']['value'],'ms/step %.2f'%d['ms_per_step'], {k:round(v['ms_per_step'],2) for k,v in d['stages'].items()})
PY
