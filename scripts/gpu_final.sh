# usage: gpu_final.sh <tag>  (run under gpurun) — smoke + pytest -m gpu + default bench + reference arm + single-frame line + matcher call latency
T=${1:-f}
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/${T}_smoke.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${T}_pytest.log
tail -3 gpurun_out/${T}_pytest.log
timeout 900 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/${T}_bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_ref.json 2>&1; echo "ref rc=$?"
timeout 300 python bench.py --workload single > gpurun_out/${T}_bench_single.json 2> gpurun_out/${T}_bench_single.err; echo "single rc=$?"; cut -c1-200 gpurun_out/${T}_bench_single.json
python scripts/exp_matcher_latency.py > gpurun_out/${T}_matcher_call_latency.json; cat gpurun_out/${T}_matcher_call_latency.json
