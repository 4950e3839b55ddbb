#!/bin/bash
# tools/gpu_multi.sh <tag> <N> — the driver's multi-GPU launch of bench.py (one rank per GPU, torchrun) + the topology of the box
tag=${1:-r02x}; N=${2:-2}
nvidia-smi topo -m > gpurun_out/${tag}_topo_${N}gpu.txt 2>&1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N \
    > gpurun_out/${tag}_bench_${N}gpu.json 2> gpurun_out/${tag}_bench_${N}gpu.err; echo "bench N=$N rc=$?"; tail -3 gpurun_out/${tag}_bench_${N}gpu.err
python - "$tag" "$N" <<'P'
import json, sys
d = json.load(open("gpurun_out/%s_bench_%sgpu.json" % (sys.argv[1], sys.argv[2])))
print("value", d["value"], "e2e", d["e2e"]["value"], "ceiling", d["e2e"].get("copy_ceiling_frames_s"), "frac", d["e2e"].get("frac_of_ceiling"))
print("cfg3", d["cfg3"].get("frames_per_s"), d["cfg3"].get("rank0_shard"), d["cfg3"].get("rank0_submit_ms"))
print("cfg5", {k: d["cfg5"].get(k) for k in ("frames_per_s", "ms_per_frame", "pairs_per_rank", "gcmp_s_e2e", "error")})
print("refset", {k: (d.get("reference_setting") or {}).get(k) for k in ("track_call_us", "batched_frames_per_s")})
P
