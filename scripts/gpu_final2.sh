# usage: gpu_final2.sh <tag>  (run under gpurun) — the round's closing evidence on the FINAL sources:
# smoke, pytest -m gpu, default bench + reference arm, latency workloads, ncu launch list, ncu --set full of one 2048-frame chunk
T=${1:-r02z}
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/${T}_smoke.log
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${T}_pytest.log
tail -3 gpurun_out/${T}_pytest.log
timeout 900 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref rc=$?"
python - $T <<'PY'
import json,sys
d=json.loads(open(f'gpurun_out/{sys.argv[1]}_bench.json').read().strip().splitlines()[-1])
print('value',round(d['value']),'e2e',round(d['e2e']['value']),'ms/step',round(d['ms_per_step'],2),'roofline',round(d['roofline']['frac'],4), 'hamming', d.get('hamming',{}).get('value'))
r=json.loads(open(f'gpurun_out/{sys.argv[1]}_bench_ref.json').read().strip().splitlines()[-1]); print('reference arm', r['value'])
PY
for w in single rgbd_single matcher_calls; do
  timeout 600 python bench.py --workload $w > gpurun_out/${T}_bench_$w.json 2> gpurun_out/${T}_bench_$w.err; echo "$w rc=$?"; cut -c1-160 gpurun_out/${T}_bench_$w.json
done
LCMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-hamming"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"pyr_|blur7|fast_|octree|describe|grid_build|search_init|repitch|knn2" --csv --log-file gpurun_out/${T}_launches.csv $LCMD > gpurun_out/${T}_launches.log 2>&1; echo "launch list rc=$?"
CMD="python bench.py --frames 2048 --chunk 2048 --steps 1 --warmup 3 --no-cpu --no-hamming"
$CMD > gpurun_out/${T}_ncu_plain.log 2>&1 || exit 1
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'pyr_|blur7|fast_|octree|describe_kernel|grid_build|search_init' -s 39 -c 13 \
    -f -o gpurun_out/${T}_chunk2048 $CMD > gpurun_out/${T}_ncu_chunk2048.log 2>&1; echo "ncu chunk2048 rc=$?"
ls -la gpurun_out/${T}*.ncu-rep
