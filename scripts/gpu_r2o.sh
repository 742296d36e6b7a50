#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_match2_gpu.py tests/test_kfmatch_gpu.py -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/r2o_pytest.log
ORBX_LAT_TRACE=1 timeout 300 python scripts/exp_matcher_latency.py > gpurun_out/r2o_lat_trace.json 2> gpurun_out/r2o_lat_trace.log
timeout 300 python scripts/exp_matcher_latency.py > gpurun_out/r2o_lat.json 2> gpurun_out/r2o_lat.err
tail -5 gpurun_out/r2o_lat_trace.log; cat gpurun_out/r2o_lat.json
