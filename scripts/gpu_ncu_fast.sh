# usage: gpu_ncu_fast.sh <tag> — ncu --set full of the FAST kernel of one 512-frame chunk (line profile: scripts/ncu_lines.py <rep> fast_cells)
T=${1:-fast}
mkdir -p gpurun_out
CMD="python bench.py --frames 512 --chunk 512 --steps 1 --warmup 3 --no-cpu --no-hamming"
$CMD > gpurun_out/${T}_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fast_cells -s 3 -c 1 -f -o gpurun_out/${T}_fast $CMD > gpurun_out/${T}_ncu.log 2>&1; echo "rc=$?"
