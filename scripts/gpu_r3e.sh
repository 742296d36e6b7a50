#!/bin/bash
T=r02z
mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'pyr_|blur7|fast_|octree|describe|grid_build|search_init|repitch|knn2' --csv --log-file gpurun_out/${T}_launches.csv \
    python bench.py --no-cpu --steps 2 --warmup 3 > gpurun_out/${T}_ncu_launch.log 2>&1; echo "launch list rc=$?"
wc -l gpurun_out/${T}_launches.csv
