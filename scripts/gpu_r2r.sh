#!/bin/bash
mkdir -p gpurun_out
ORBX_LAT_TRACE=1 timeout 600 python bench.py --workload matcher_calls --steps 2 --no-cpu > gpurun_out/r2r_trace.json 2> gpurun_out/r2r_trace.log
grep "orbx lat" gpurun_out/r2r_trace.log | sort | uniq -c | sort -k3 | awk '{$1="";print}' | sort -u -k3,3
