#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scripts/exp_call_once.py init > gpurun_out/r2w_plain.log 2>&1 || { tail -5 gpurun_out/r2w_plain.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:search_init_kernel -s 3 -c 1 -f -o gpurun_out/r2w_init python scripts/exp_call_once.py init > gpurun_out/r2w_ncu.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:search_bow_kernel -s 3 -c 1 -f -o gpurun_out/r2w_bow python scripts/exp_call_once.py bow >> gpurun_out/r2w_ncu.log 2>&1
grep -E "PROF|rror" gpurun_out/r2w_ncu.log | tail -6; ls -la gpurun_out/r2w_*
