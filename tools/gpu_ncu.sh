#!/bin/bash
# tools/gpu_ncu.sh <tag> — the two ncu passes of B200_PROFILING.md over the resident step (tools/ncu_step.py), each after the plain
# run of the same command has exited 0.  Graph replay off so that every kernel is an ordinary launch (same kernels either way).
tag=${1:-r02x}
export KFS_ORB_GRAPHS=0
python tools/ncu_step.py 4 > gpurun_out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 14 -c 60 --csv --log-file gpurun_out/${tag}_launches.csv \
    python tools/ncu_step.py 4 > gpurun_out/${tag}_ncu1.log 2>&1
echo "launch list rc=$?"
python tools/ncu_step.py 2 > gpurun_out/${tag}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_ -s 27 -c 14 -f -o gpurun_out/${tag}_prof \
    python tools/ncu_step.py 2 > gpurun_out/${tag}_ncu2.log 2>&1
echo "full capture rc=$?"; tail -3 gpurun_out/${tag}_ncu2.log; ls -la gpurun_out/
