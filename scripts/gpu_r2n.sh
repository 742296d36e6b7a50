#!/bin/bash
# phase breakdown of the stateless matcher calls (ORBX_LAT_TRACE) next to the call latency
mkdir -p gpurun_out
ORBX_LAT_TRACE=1 timeout 300 python scripts/exp_matcher_latency.py > gpurun_out/r2n_lat_trace.json 2> gpurun_out/r2n_lat_trace.log
timeout 300 python scripts/exp_matcher_latency.py > gpurun_out/r2n_lat.json 2> gpurun_out/r2n_lat.err
tail -20 gpurun_out/r2n_lat_trace.log; cat gpurun_out/r2n_lat.json
