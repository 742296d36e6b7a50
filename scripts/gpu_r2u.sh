#!/bin/bash
mkdir -p gpurun_out
echo "== alone"; timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r2u_noalloc_alone.log
echo "== whole suite"; timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | tee gpurun_out/r2u_pytest_all.log
