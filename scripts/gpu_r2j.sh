# usage: gpu_r2j.sh <tag> <N> — N-GPU: topology, H2D sharing table, sharded kNN test, torchrun bench at N
T=${1:-r02j}; N=${2:-8}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/${T}_topo_${N}.txt 2>&1
timeout 300 python scripts/exp_h2d_topology.py > gpurun_out/${T}_h2d_topology_${N}.json 2> gpurun_out/${T}_h2d_topology.err; echo "h2d rc=$?"
cat gpurun_out/${T}_h2d_topology_${N}.json | cut -c1-600
timeout 600 python -m pytest tests/test_sharded_gpu.py -m gpu -x -q > gpurun_out/${T}_pytest_sharded_${N}.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/${T}_pytest_sharded_${N}.log
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,COLL timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 \
    bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/${T}_bench_${N}.json 2> gpurun_out/${T}_bench_${N}.err; echo "bench rc=$?"
grep -c "AllGather" gpurun_out/${T}_bench_${N}.err; grep -m2 "AllGather" gpurun_out/${T}_bench_${N}.err | cut -c1-200 > gpurun_out/${T}_comm_log_tail_${N}.txt
python - $T $N <<'PY'
import json,sys
txt=open(f'gpurun_out/{sys.argv[1]}_bench_{sys.argv[2]}.json').read().strip().splitlines()
print('stdout lines', len(txt))
d=json.loads(txt[-1])
print('value',round(d['value']),'e2e',round(d['e2e']['value']),'ms/step',round(d['ms_per_step'],2), 'h2d', d['e2e'].get('h2d_gbs_measured'))
h=d['hamming']; print({k:h[k] for k in ('value','per_gpu','nranks','ms','exchange_ms_alone','local_scan_cmp_s_per_gpu','efficiency_vs_n1')})
PY
