# usage: gpu_multi_both_arms.sh <tag> <N> — N-GPU: sharded kNN test + torchrun bench at N (both arms)
T=${1:-r02z}; N=${2:-8}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_sharded_gpu.py -m gpu -x -q > gpurun_out/${T}_pytest_sharded_${N}.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/${T}_pytest_sharded_${N}.log
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,COLL timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 \
    bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/${T}_bench_${N}gpu.json 2> gpurun_out/${T}_bench_${N}gpu.err; echo "bench rc=$?"
grep -c "AllGather" gpurun_out/${T}_bench_${N}gpu.err; grep -m2 "AllGather" gpurun_out/${T}_bench_${N}gpu.err | cut -c1-200 > gpurun_out/${T}_comm_log_tail_${N}gpu.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29612 \
    bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/${T}_bench_ref_${N}gpu.json 2> gpurun_out/${T}_bench_ref_${N}gpu.err; echo "ref rc=$?"
python - $T $N <<'PY'
import json,sys
txt=open(f'gpurun_out/{sys.argv[1]}_bench_{sys.argv[2]}gpu.json').read().strip().splitlines()
print('stdout lines', len(txt))
d=json.loads(txt[-1])
print('value',round(d['value']),'e2e',round(d['e2e']['value']),'ms/step',round(d['ms_per_step'],2), 'h2d', d['e2e'].get('h2d_gbs_measured'))
h=d['hamming']; print({k:h[k] for k in ('value','per_gpu','nranks','ms','exchange_ms_alone','local_scan_cmp_s_per_gpu','efficiency_vs_n1')})
r=open(f'gpurun_out/{sys.argv[1]}_bench_ref_{sys.argv[2]}gpu.json').read().strip().splitlines(); print('ref lines', len(r), r[-1][:200])
PY
