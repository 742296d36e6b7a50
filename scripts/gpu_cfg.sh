# extractor parity tests + hd / stereo / frames workloads (run under gpurun)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_extract_gpu.py tests/test_match_gpu.py tests/test_match2_gpu.py -m gpu -x -q 2>&1 | tail -3
for wl in ${@:-hd stereo frames}; do
  timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu > gpurun_out/cfg_$wl.json 2> gpurun_out/cfg_$wl.err; echo "$wl rc=$?"
  python - $wl <<'PY'
import json,sys
d=json.loads(open(f'gpurun_out/cfg_{sys.argv[1]}.json').read().strip().splitlines()[-1])
print(sys.argv[1], 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'ms/step %.2f'%d['ms_per_step'], {k:round(v['ms_per_step'],2) for k,v in d.get('stages',{}).items()})
PY
done
