# usage: gpu_r2k.sh <tag> — voxel + knn + sharded tests, knn bench, default-bench hamming leg
T=${1:-r02k}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_voxel.py tests/test_match2_gpu.py tests/test_sharded_gpu.py tests/test_fullsize_gpu.py -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/${T}_pytest.log
for Q in 65536 262144 1048576; do
timeout 300 python bench.py --workload knn --queries $Q --steps 2 --warmup 1 > gpurun_out/${T}_knn_$Q.json 2> gpurun_out/${T}_knn_$Q.err
python -c "
import json; d=json.loads(open('gpurun_out/${T}_knn_$Q.json').read().strip().splitlines()[-1]); print($Q, 'cmp/s', d['value'], 'frac', d['roofline']['frac'], 'peak', d['roofline']['peak'])"
done
