#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_match2_gpu.py -m gpu -x -q 2>&1 | tail -4
timeout 1200 python scripts/soak_parity.py --trials 6000 --seed 7 --budget-s 600 > gpurun_out/r02z_soak.log 2>&1; echo "soak rc=$?" >> gpurun_out/r02z_soak.log
tail -4 gpurun_out/r02z_soak.log
