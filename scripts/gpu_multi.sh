# multi-GPU check (run under gpurun --gpus N): frames + knn under torchrun
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L | head -8
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/multi_frames_$N.json 2> gpurun_out/multi_frames_$N.err; echo "frames rc=$?"; tail -3 gpurun_out/multi_frames_$N.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 3 --warmup 2 --workload knn > gpurun_out/multi_knn_$N.json 2> gpurun_out/multi_knn_$N.err; echo "knn rc=$?"; tail -3 gpurun_out/multi_knn_$N.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 3 --warmup 2 --workload stereo > gpurun_out/multi_stereo_$N.json 2> gpurun_out/multi_stereo_$N.err; echo "stereo rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus $N --impl reference --steps 1 --warmup 1 > gpurun_out/multi_ref_$N.json 2>&1; echo "ref rc=$?"
python - $N <<'PY'
import json,sys
for wl in ('frames','knn','stereo','ref'):
    try:
        d=json.loads([l for l in open(f'gpurun_out/multi_{wl}_{sys.argv[1]}.json').read().splitlines() if l.startswith('{')][-1])
        print(wl, 'n_gpus',d['n_gpus'],'value %.4g'%d['value'],'e2e %.4g'%d['e2e']['value'],'ms/step %.2f'%d['ms_per_step'])
    except Exception as e: print(wl,'parse fail',e)
PY
