#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --frames 512 --chunk 512 --steps 1 --warmup 3 --no-cpu --no-hamming"
$CMD > gpurun_out/ncu_plain.log 2>&1 || { tail -3 gpurun_out/ncu_plain.log; exit 1; }
# the third warm-up step's kernels: skip the first 2 steps x 13 launches
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"pyr_resize4|blur7|describe" -s 20 -c 9 -f -o gpurun_out/r3d_kernels $CMD > gpurun_out/ncu_run.log 2>&1
echo "rc=$?"; grep -E "PROF|rror" gpurun_out/ncu_run.log | tail -3
