# usage: gpu_soaks.sh <tag> [extractor budget s] [matcher budget s] — both randomised parity soaks on the current build
T=${1:-soak}; mkdir -p gpurun_out
timeout 1200 python scripts/soak_parity.py --trials 6000 --seed 7 --budget-s ${2:-420} > gpurun_out/${T}_soak.log 2>&1; echo "soak rc=$?" >> gpurun_out/${T}_soak.log; tail -2 gpurun_out/${T}_soak.log
timeout 1200 python scripts/soak_matchers.py --trials 30000 --seed 3 --budget-s ${3:-240} > gpurun_out/${T}_soak_matchers.log 2>&1; echo "soak rc=$?" >> gpurun_out/${T}_soak_matchers.log; tail -2 gpurun_out/${T}_soak_matchers.log
