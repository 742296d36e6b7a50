# usage: gpu_single.sh <tag>  (run under gpurun) — pytest -m gpu + the single-frame latency line (graph replay and plain launches)
T=${1:-s}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${T}_pytest.log
tail -5 gpurun_out/${T}_pytest.log
timeout 600 python bench.py --workload single > gpurun_out/${T}_bench_single.json 2> gpurun_out/${T}_bench_single.err; echo "single rc=$?"
tail -3 gpurun_out/${T}_bench_single.err; cat gpurun_out/${T}_bench_single.json
ORBX_NO_GRAPH=1 timeout 600 python bench.py --workload single --no-cpu > gpurun_out/${T}_bench_single_nograph.json 2> gpurun_out/${T}_bench_single_nograph.err; echo "nograph rc=$?"
cat gpurun_out/${T}_bench_single_nograph.json
