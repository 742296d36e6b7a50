# randomised parity soak under gpurun: bash scripts/gpu_soak.sh <tag> [trials] [seed]
mkdir -p gpurun_out
TAG=${1:-soak}; TRIALS=${2:-400}; SEED=${3:-1}
timeout 900 python scripts/soak_parity.py --trials $TRIALS --seed $SEED --budget-s 420 > gpurun_out/${TAG}_soak.log 2>&1; echo "soak rc=$?" >> gpurun_out/${TAG}_soak.log
tail -4 gpurun_out/${TAG}_soak.log
