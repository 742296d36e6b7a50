#!/bin/bash
mkdir -p gpurun_out
timeout 900 python scripts/soak_matchers.py --trials ${1:-300} --seed ${2:-1} --budget-s ${3:-240} > gpurun_out/r02z_soak_matchers.log 2>&1; echo "soak rc=$?" >> gpurun_out/r02z_soak_matchers.log
tail -6 gpurun_out/r02z_soak_matchers.log
