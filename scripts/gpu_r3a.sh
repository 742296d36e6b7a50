#!/bin/bash
mkdir -p gpurun_out
for ch in 512 256; do echo "== chunk $ch"; timeout 300 python scripts/exp_pipe_trace.py $ch 2>&1 | grep -E "^chunk|total|pipe" | tail -24; done
