# usage: gpu_r2f.sh <tag> — all GPU tests, knn bench (packed-key top-2), rgbd_single latency, default bench
T=${1:-r02f}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${T}_pytest.log
tail -6 gpurun_out/${T}_pytest.log
timeout 600 python bench.py --workload rgbd_single > gpurun_out/${T}_rgbd_single.json 2> gpurun_out/${T}_rgbd_single.err; echo "rgbd_single rc=$?"
tail -2 gpurun_out/${T}_rgbd_single.err; python -c "
import json; d=json.loads(open('gpurun_out/${T}_rgbd_single.json').read().strip().splitlines()[-1]); print(d['latency_us'], d.get('cpu_baseline'))"
timeout 600 python bench.py --workload single > gpurun_out/${T}_single.json 2> gpurun_out/${T}_single.err; echo "single rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/${T}_single.json').read().strip().splitlines()[-1]); print(d['latency_us'])"
timeout 900 python bench.py --no-cpu > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
tail -3 gpurun_out/${T}_bench.err
python - $T <<'PY'
import json,sys
d=json.loads(open(f'gpurun_out/{sys.argv[1]}_bench.json').read().strip().splitlines()[-1])
print('value',round(d['value']),'e2e',round(d['e2e']['value']),'ms/step',round(d['ms_per_step'],2))
h=d['hamming']; print({k:h[k] for k in ('value','ms')}, h['popc_roofline'])
print(d['roofline'])
PY
