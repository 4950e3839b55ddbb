#!/bin/bash
# compute-sanitizer over the product kernels (VERDICT r2 "missing 2"): memcheck, racecheck, synccheck, initcheck on a reduced
# selection of the GPU suite, CUDA graphs off so that every kernel is a plain launch the tools can attribute.
#   gpurun --timeout 3000 -- 'bash tools/gpu_sanitize.sh r03'     -> gpurun_out/<tag>_sanitizer_<tool>.log (+ _summary.txt)
# Each tool is bounded by its own timeout; a tool that runs out of time says so in the summary instead of hanging the box.
# NOTE (round 3): compute-sanitizer is closed on the GPU pool this repository is measured on (profiles/r03b_sanitizer_closed.txt);
# the script is kept for a box where the tool runs.
TAG=${1:-r03}
TOOLS=${2:-"memcheck synccheck initcheck racecheck"}
PER_TOOL=${3:-900}
OUT=gpurun_out
mkdir -p $OUT
export KFS_ORB_GRAPHS=0
SEL_ORB="tests/test_gpu_parity.py::test_orb_matches_cv2_golden[cfg1_480p_a] tests/test_gpu_parity.py::test_orb_matches_cv2_golden[odd_333x257] tests/test_gpu_parity.py::test_orb_matches_cv2_golden[cfg1_480p_masked]"
SEL_KNN="tests/test_gpu_parity.py::test_knn_vs_oracle_shapes tests/test_gpu_switches.py::test_knn_cluster_splits_with_tiny_device_side_train_count tests/test_zz_no_alloc.py"
SEL_SW="tests/test_gpu_switches.py::test_switches_do_not_change_results tests/test_gpu_switches.py::test_knn_split_and_cluster_merge_do_not_change_results"
SEL_RT="tests/test_gpu_replay.py::test_replay_keyframe_cache_across_batches tests/test_gpu_track.py::test_track_one_call_per_frame"
: > $OUT/${TAG}_sanitizer_summary.txt
for tool in $TOOLS; do
    log=$OUT/${TAG}_sanitizer_${tool}.log
    extra=""
    [ "$tool" = "memcheck" ] && extra="--leak-check no"
    [ "$tool" = "initcheck" ] && extra="--track-unused-memory no"
    sel="$SEL_ORB $SEL_KNN $SEL_SW $SEL_RT"
    # racecheck serialises shared-memory accesses of every warp: the switch sweeps (40 handles x 5 frames) would not finish
    [ "$tool" = "racecheck" ] && sel="$SEL_ORB $SEL_KNN $SEL_RT"
    t0=$(date +%s)
    timeout $PER_TOOL compute-sanitizer --tool $tool $extra --error-exitcode 9 --print-limit 40 \
        python -m pytest $sel -x -q -m gpu -p no:cacheprovider > $log 2>&1
    rc=$?
    t1=$(date +%s)
    echo "$tool rc=$rc seconds=$((t1 - t0)) :: $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $log | tail -1) :: $(grep -E ' passed| failed| error' $log | tail -1)" | tee -a $OUT/${TAG}_sanitizer_summary.txt
done
