# usage: gpu_round.sh <tag>  (run under gpurun) — smoke + pytest -m gpu + default bench + ncu launch list + one ncu --set full
T=${1:-g}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/${T}_smi.txt 2>&1; nproc >> gpurun_out/${T}_smi.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/${T}_smoke.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${T}_pytest.log
tail -3 gpurun_out/${T}_pytest.log
timeout 900 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
tail -2 gpurun_out/${T}_bench.err
python - $T <<'PY'
import json,sys
d=json.loads(open(f'gpurun_out/{sys.argv[1]}_bench.json').read().strip().splitlines()[-1])
print('value',round(d['value']),'e2e',round(d['e2e']['value']),'ms/step',round(d['ms_per_step'],2))
for k,v in d['stages'].items(): print(f"  {k:12s} {v['ms_per_step']:8.3f} ms  {v['launches_per_step']:.0f} launches")
print(d['roofline']); print(d.get('cpu_baseline')); print(d.get('clocks'))
PY
# launch list of the same command (cold-cache, serialised: shares must agree, not absolutes)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'pyr_|blur7|fast_|octree|describe|grid_build|search_init|repitch' --csv --log-file gpurun_out/${T}_launches.csv \
    python bench.py --no-cpu --steps 2 --warmup 3 > gpurun_out/${T}_ncu_launch.log 2>&1; echo "launch list rc=$?"
# one --set full capture of every extractor/matcher kernel of one 64-frame chunk
CMD="python bench.py --frames 64 --chunk 64 --steps 1 --warmup 3 --no-cpu"
$CMD > gpurun_out/${T}_ncu_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on \
    -k regex:'pyr_|blur7|fast_|octree|describe_kernel|grid_build|search_init' -s 39 -c 16 \
    -f -o gpurun_out/${T}_full $CMD > gpurun_out/${T}_ncu_full.log 2>&1
echo "ncu full rc=$?"; ls -la gpurun_out/*.ncu-rep
