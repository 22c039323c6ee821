#!/bin/sh
# compute-sanitizer pass over smoke() and one parity test per kernel family (run on the GPU box):
#   sh tools/sanitize.sh memcheck|racecheck|initcheck|synccheck  -> gpurun_out/r2_sanitizer_<tool>.log
# The 4K / 40000-feature cases are left out (hours under the sanitizer); everything else of the -m gpu suite that
# launches a distinct kernel or template variant is in.
TOOL="${1:-memcheck}"
OUT="gpurun_out/r2_sanitizer_${TOOL}.log"
mkdir -p gpurun_out
SEL='smoke or (test_extract_stages_match_oracle and (small or C1 or C3-fisheye-partial)) or test_batch_equals_single or test_flat_and_low_texture or test_clustered_keypoints_deep_quadtree or test_single_frame_path_state_changes or test_equal_count_ties or test_stereo_euroc or test_stereo_single_frame_api or test_extract_stereo_one_call or test_knn2 or test_search_by_projection or test_window_candidates or test_search_for_initialization or test_search_for_triangulation or test_search_by_bow or test_search_by_sim3 or test_fuse or test_distinctive or test_bow or test_undistort or test_search_local_points or test_prepare or test_stereo_fisheye or test_noalloc'
compute-sanitizer --tool "$TOOL" --error-exitcode 9 --print-limit 30 --target-processes all \
    python -m pytest tests -q -m gpu -x -k "$SEL" -p no:cacheprovider > "$OUT" 2>&1
echo "exit=$?" >> "$OUT"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|exit=" "$OUT" | tail -8
