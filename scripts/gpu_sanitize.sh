# usage: gpu_sanitize.sh <tag>  (1 GPU) — compute-sanitizer memcheck / racecheck / initcheck / synccheck over one small case per kernel family
T=${1:-r02s}
mkdir -p gpurun_out
MEM="tests/test_extract_gpu.py::test_tall_and_wide_images tests/test_extract_gpu.py::test_octree_global_key_path tests/test_extract_gpu.py::test_single_frame_graph_replay \
 tests/test_extract_gpu.py::test_wide_image_on_larger_handle_grows_capacity tests/test_match_gpu.py tests/test_match2_gpu.py::test_search_by_projection \
 tests/test_match2_gpu.py::test_compute_stereo_matches_stateless tests/test_match2_gpu.py::test_extract_stereo_batch_padded_host_rows \
 tests/test_match2_gpu.py::test_knn2_sharded_merge_equals_whole tests/test_sharded_gpu.py::test_sharded_entry_point_one_rank_walks_the_exchange \
 tests/test_bow_gpu.py tests/test_kfmatch_gpu.py tests/test_frame_gpu.py"
RACE="tests/test_extract_gpu.py::test_tall_and_wide_images tests/test_extract_gpu.py::test_octree_global_key_path tests/test_match_gpu.py::test_search_init_collisions_and_duplicates \
 tests/test_match2_gpu.py::test_search_by_projection tests/test_match2_gpu.py::test_compute_stereo_matches_stateless tests/test_bow_gpu.py tests/test_kfmatch_gpu.py"
run() { # tool, timeout, tests
  tool=$1; to=$2; shift 2
  timeout $to compute-sanitizer --tool $tool --print-limit 20 --error-exitcode 9 python -m pytest $@ -m gpu -x -q -p no:cacheprovider > gpurun_out/${T}_$tool.log 2>&1
  echo "$tool rc=$?" | tee -a gpurun_out/${T}_$tool.log
  grep -E "ERROR SUMMARY|passed|failed|RACECHECK SUMMARY" gpurun_out/${T}_$tool.log | tail -4
}
run memcheck 1500 $MEM
run racecheck 1500 $RACE
run initcheck 1200 $RACE
run synccheck 900 $RACE
