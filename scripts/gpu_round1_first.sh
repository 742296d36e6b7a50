set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r1_smi.txt 2>&1
nproc >> gpurun_out/r1_smi.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/r1_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r1_smoke.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r1_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1_pytest.log
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/r1_bench.json 2> gpurun_out/r1_bench.err; echo "bench rc=$?" >> gpurun_out/r1_bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1_bench_ref.json 2> gpurun_out/r1_bench_ref.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1_launches.csv python bench.py --steps 1 --warmup 3 --frames 512 --no-cpu > gpurun_out/r1_ncu.log 2>&1
tail -3 gpurun_out/r1_smoke.log gpurun_out/r1_pytest.log gpurun_out/r1_bench.err; cat gpurun_out/r1_bench.json
