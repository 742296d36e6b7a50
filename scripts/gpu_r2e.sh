# usage: gpu_r2e.sh <tag> — bench with the CPU arms, ncu --set full: whole 512-frame chunk (ncu_constants) + matcher family pipe metrics
T=${1:-r02e}
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref rc=$?"
python - $T <<'PY'
import json,sys
d=json.loads(open(f'gpurun_out/{sys.argv[1]}_bench.json').read().strip().splitlines()[-1])
print('value',round(d['value']),'e2e',round(d['e2e']['value']),'ms/step',round(d['ms_per_step'],2)); print(json.dumps(d['cpu_baseline'],indent=1))
r=json.loads(open(f'gpurun_out/{sys.argv[1]}_bench_ref.json').read().strip().splitlines()[-1]); print('reference arm', r['value'], r['config']['implementation'])
PY
CMD="python bench.py --frames 512 --chunk 512 --steps 1 --warmup 3 --no-cpu --no-hamming"
$CMD > gpurun_out/${T}_ncu_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'pyr_|blur7|fast_|octree|describe_kernel|grid_build|search_init' -s 39 -c 13 \
    -f -o gpurun_out/${T}_chunk512 $CMD > gpurun_out/${T}_ncu_chunk512.log 2>&1; echo "ncu chunk512 rc=$?"
KCMD="python bench.py --workload knn --queries 131072 --db-per-gpu 1000000 --steps 1 --warmup 1"
$KCMD > gpurun_out/${T}_knn_plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:knn2_kernel -s 1 -c 1 \
    -f -o gpurun_out/${T}_knn2 $KCMD > gpurun_out/${T}_ncu_knn2.log 2>&1; echo "ncu knn2 rc=$?"
SCMD="python bench.py --workload stereo --steps 1 --warmup 1"
$SCMD > gpurun_out/${T}_stereo_plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:'stereo_' -s 2 -c 2 \
    -f -o gpurun_out/${T}_stereo $SCMD > gpurun_out/${T}_ncu_stereo.log 2>&1; echo "ncu stereo rc=$?"
PCMD="python scripts/exp_matcher_latency.py"
$PCMD > gpurun_out/${T}_matcher_latency.json 2> gpurun_out/${T}_matcher_latency.err && timeout 600 ncu --set full --clock-control none --import-source on -k regex:'search_proj' -s 10 -c 2 \
    -f -o gpurun_out/${T}_proj $PCMD > gpurun_out/${T}_ncu_proj.log 2>&1; echo "ncu proj rc=$?"
cat gpurun_out/${T}_matcher_latency.json
ls -la gpurun_out/${T}*.ncu-rep
