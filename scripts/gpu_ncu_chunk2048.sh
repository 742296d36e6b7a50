# usage: gpu_ncu_chunk2048.sh <tag> — ncu --set full of every extractor / matcher kernel of one 2048-frame chunk (the chunk bench.py launches);
# feed the report to scripts/ncu_constants.py <rep> 2048 <label> to refresh profiles/ncu_constants.json
T=${1:-r02z}
mkdir -p gpurun_out
CMD="python bench.py --frames 2048 --chunk 2048 --steps 1 --warmup 3 --no-cpu --no-hamming"
$CMD > gpurun_out/${T}_ncu_plain.log 2>&1 || exit 1
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'pyr_|blur7|fast_|octree|describe_kernel|grid_build|search_init' -s 39 -c 13 \
    -f -o gpurun_out/${T}_chunk2048 $CMD > gpurun_out/${T}_ncu_chunk2048.log 2>&1; echo "ncu chunk2048 rc=$?"
ls -la gpurun_out/${T}_chunk2048.ncu-rep
