# usage: gpu_r2g.sh <tag> — extractor / frame tests + the two latency lines
T=${1:-r02g}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_extract_gpu.py tests/test_frame_gpu.py tests/test_host_api.py -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${T}_pytest.log
tail -6 gpurun_out/${T}_pytest.log
timeout 600 python bench.py --workload rgbd_single --no-cpu > gpurun_out/${T}_rgbd_single.json 2> gpurun_out/${T}_rgbd_single.err; echo "rgbd_single rc=$?"
tail -2 gpurun_out/${T}_rgbd_single.err; python -c "
import json; d=json.loads(open('gpurun_out/${T}_rgbd_single.json').read().strip().splitlines()[-1]); print(d['latency_us'])"
timeout 600 python bench.py --workload single --no-cpu > gpurun_out/${T}_single.json 2> gpurun_out/${T}_single.err; echo "single rc=$?"
tail -2 gpurun_out/${T}_single.err; python -c "
import json; d=json.loads(open('gpurun_out/${T}_single.json').read().strip().splitlines()[-1]); print(d['latency_us'])"
