# all bench workloads at N=1 (run under gpurun)
mkdir -p gpurun_out
for wl in frames hd stereo knn; do
  timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; echo "$wl rc=$?"
  tail -2 gpurun_out/bench_$wl.err
  python - $wl <<'PY'
import json,sys
try:
    d=json.loads(open(f'gpurun_out/bench_{sys.argv[1]}.json').read().strip().splitlines()[-1])
    print(sys.argv[1], d['metric'], 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'ms/step %.2f'%d['ms_per_step'], 'roof', round(d['roofline']['frac'],4))
    if 'stages' in d: print('   ', {k:round(v['ms_per_step'],2) for k,v in d['stages'].items()})
    if 'hamming' in d: print('    hamming', d['hamming'])
    if 'cpu_baseline' in d: print('    cpu', d['cpu_baseline'])
except Exception as e: print('parse fail', e)
PY
done
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2>&1; tail -c 600 gpurun_out/bench_ref.json
