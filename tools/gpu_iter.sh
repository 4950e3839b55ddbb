#!/bin/bash
# tools/gpu_iter.sh <tag> [ncu] — one development round trip: the GPU suite, a short bench with stage times, and (with "ncu")
# the two ncu passes of tools/gpu_ncu.sh — those only when the suite is green.
tag=${1:-r02x}
python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; rc=$?; echo "pytest rc=$rc"; tail -4 gpurun_out/${tag}_pytest.log
python bench.py --no-cpu-baseline --steps 10 --e2e-seconds 0.3 --sustain 0 --cfg3-frames 2000 --cfg5-frames 4 > gpurun_out/${tag}_q.json 2> gpurun_out/${tag}_q.err || tail -5 gpurun_out/${tag}_q.err
python - "$tag" <<'P'
import json, sys
d = json.load(open("gpurun_out/%s_q.json" % sys.argv[1]))
print("fps", d["value"], "e2e", d["e2e"]["value"], " ".join("%s=%.4f" % (k["stage"], k["ms_per_batch"]) for k in d["kernels"]), "knn=%.4f" % d["roofline_popc"]["ms_per_batch"])
print("latency", {k: v for k, v in d["stream_latency"].items() if k.endswith("_us")})
print("cfg3", d["cfg3"].get("frames_per_s"), d["cfg3"].get("rank0_submit_ms"))
print("window", d["window_matching"]["one_frame"], d["window_matching"]["batch"])
P
if [ "$2" = "ncu" ] && [ $rc -eq 0 ]; then bash tools/gpu_ncu.sh $tag; fi
