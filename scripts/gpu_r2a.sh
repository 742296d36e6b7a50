# usage: gpu_r2a.sh <tag>  (under gpurun) — smoke + pytest -m gpu + short bench + ncu --set full of the FAST kernel on the 512-frame chunk
T=${1:-r02a}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/${T}_smi.txt 2>&1; nproc >> gpurun_out/${T}_smi.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/${T}_smoke.log
tail -3 gpurun_out/${T}_smoke.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${T}_pytest.log
tail -15 gpurun_out/${T}_pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
tail -2 gpurun_out/${T}_bench.err
python - $T <<'PY'
import json,sys
d=json.loads(open(f'gpurun_out/{sys.argv[1]}_bench.json').read().strip().splitlines()[-1])
print('value',round(d['value']),'e2e',round(d['e2e']['value']),'ms/step',round(d['ms_per_step'],2))
for k,v in d['stages'].items(): print(f"  {k:12s} {v['ms_per_step']:8.3f} ms  {v['launches_per_step']:.0f} launches")
print(d['roofline'])
PY
CMD="python bench.py --frames 512 --chunk 512 --steps 1 --warmup 3 --no-cpu"
$CMD > gpurun_out/${T}_ncu_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'fast_' -s 3 -c 1 \
    -f -o gpurun_out/${T}_fast512 $CMD > gpurun_out/${T}_ncu_fast.log 2>&1
echo "ncu rc=$?"; grep -E "PROF|rror" gpurun_out/${T}_ncu_fast.log | tail -3; ls -la gpurun_out/*.ncu-rep
