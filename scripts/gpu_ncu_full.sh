# one ncu --set full capture of every kernel of one 64-frame chunk (run under gpurun)
mkdir -p gpurun_out
CMD="python bench.py --frames 64 --chunk 64 --steps 1 --warmup 3 --no-cpu"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on \
    -k regex:'pyr_|blur7|fast_cells|octree_kernel|describe_kernel|grid_build|search_init' -s 39 -c 16 \
    -f -o gpurun_out/${1:-prof} $CMD > gpurun_out/ncu_run.log 2>&1
echo "rc=$?"; tail -5 gpurun_out/ncu_run.log; ls -la gpurun_out/*.ncu-rep
