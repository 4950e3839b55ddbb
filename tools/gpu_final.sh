#!/bin/bash
# tools/gpu_final.sh <tag> — the round's record: GPU suite, full default bench (both arms), ncu launch list + full capture, k_describe capture
tag=${1:-r02z}
bash tools/gpu_session.sh $tag
bash tools/gpu_ncu.sh $tag
