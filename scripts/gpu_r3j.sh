#!/bin/bash
timeout 900 python -m pytest tests/test_sor.py tests/test_voxel.py -m gpu -x -q 2>&1 | tail -6
timeout 600 python scripts/exp_sor_timing.py 2>&1 | tail -5
