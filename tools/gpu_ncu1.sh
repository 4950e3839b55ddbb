#!/bin/bash
# tools/gpu_ncu1.sh <tag> — ncu launch list of the one-frame-per-call chain (8 frames, one per call: 7 extraction launches + the
# single-pair kNN each) and a --set full capture of k_describe2 / the cluster kNN; graphs off so every kernel is a plain launch
tag=${1:-r02x}
export KFS_ORB_GRAPHS=0
python tools/ncu_step.py 2 8 1 > gpurun_out/${tag}_one_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 16 -c 64 --csv --log-file gpurun_out/${tag}_launches_one_frame.csv \
    python tools/ncu_step.py 2 8 1 > gpurun_out/${tag}_one_ncu.log 2>&1
echo "one-frame launch list rc=$?"
python tools/ncu_step.py 2 > gpurun_out/${tag}_plain3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_describe -s 1 -c 1 -f -o gpurun_out/${tag}_prof_describe \
    python tools/ncu_step.py 2 > gpurun_out/${tag}_ncu3.log 2>&1
echo "describe capture rc=$?"
