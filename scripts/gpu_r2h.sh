# usage: gpu_r2h.sh <tag> — frame tests + rgbd_single latency with 1/2/4/8 colour slices
T=${1:-r02h}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_frame_gpu.py -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${T}_pytest.log
for S in 1 2 4 8; do
ORBX_RGBD_SLICES=$S timeout 300 python bench.py --workload rgbd_single --no-cpu > gpurun_out/${T}_rgbd_single_$S.json 2> gpurun_out/${T}_rgbd_single_$S.err
python -c "
import json; d=json.loads(open('gpurun_out/${T}_rgbd_single_$S.json').read().strip().splitlines()[-1]); print('slices $S', d['latency_us'])"
done
