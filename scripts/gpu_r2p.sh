#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py --workload matcher_calls --steps 2 > gpurun_out/r2p_matcher_calls.json 2> gpurun_out/r2p_matcher_calls.err
tail -5 gpurun_out/r2p_matcher_calls.err; cat gpurun_out/r2p_matcher_calls.json
