# usage: gpu_r2b.sh <tag>  — extractor tests + short bench + ncu --set full (source) of FAST, resize (level 1), blur, describe on the 512-frame chunk
T=${1:-r02b}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_extract_gpu.py tests/test_match_gpu.py tests/test_match2_gpu.py -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${T}_pytest.log
tail -4 gpurun_out/${T}_pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
tail -2 gpurun_out/${T}_bench.err
python - $T <<'PY'
import json,sys
d=json.loads(open(f'gpurun_out/{sys.argv[1]}_bench.json').read().strip().splitlines()[-1])
print('value',round(d['value']),'e2e',round(d['e2e']['value']),'ms/step',round(d['ms_per_step'],2))
for k,v in d['stages'].items(): print(f"  {k:12s} {v['ms_per_step']:8.3f} ms  {v['launches_per_step']:.0f} launches")
PY
CMD="python bench.py --frames 512 --chunk 512 --steps 1 --warmup 3 --no-cpu"
$CMD > gpurun_out/${T}_ncu_plain.log 2>&1 || exit 1
for K in fast_ pyr_resize4 blur7 describe_kernel; do
timeout 600 ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 1 \
    -f -o gpurun_out/${T}_$K $CMD > gpurun_out/${T}_ncu_$K.log 2>&1
echo "ncu $K rc=$?"
done
ls -la gpurun_out/*.ncu-rep
