# usage: gpu_ncu_kernel.sh <regex> <outname> [skip] [count]  — ncu --set full of selected kernels of one 64-frame chunk
mkdir -p gpurun_out
CMD="python bench.py --frames 64 --chunk 64 --steps 1 --warmup 3 --no-cpu"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"$1" -s ${3:-3} -c ${4:-1} \
    -f -o gpurun_out/$2 $CMD > gpurun_out/ncu_run.log 2>&1
echo "rc=$?"; grep -E "PROF|rror" gpurun_out/ncu_run.log | tail -5
