#!/bin/bash
# usage: gpu_r2x.sh <which: init|bow> <kernel regex> <out name>
mkdir -p gpurun_out
timeout 300 python scripts/exp_call_once.py $1 > gpurun_out/r2x_plain.log 2>&1 || { tail -5 gpurun_out/r2x_plain.log; exit 1; }
timeout 600 ncu --set full --sampling-interval 0 --clock-control none --import-source on -k regex:$2 -s 3 -c 1 -f -o gpurun_out/$3 python scripts/exp_call_once.py $1 > gpurun_out/r2x_ncu.log 2>&1
grep -E "PROF|rror" gpurun_out/r2x_ncu.log | tail -3
