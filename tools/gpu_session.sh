#!/bin/bash
# tools/gpu_session.sh <tag> — one GPU round trip: the GPU test-suite, the pipe microbenchmark, the bench (both arms).
# Run through gpurun; everything lands in gpurun_out/<tag>_*.
tag=${1:-r02x}
python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/${tag}_pytest.log
if [ -x tools/lab/pipe_lab ]; then ./tools/lab/pipe_lab > gpurun_out/${tag}_pipe_lab.txt 2>&1; cat gpurun_out/${tag}_pipe_lab.txt; fi
python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/${tag}_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_ref.json 2> gpurun_out/${tag}_ref.err; echo "ref rc=$?"
