#!/bin/bash
# one short GPU round trip: parity tests of the extraction / matching kernels + a short bench with stage times
python -m pytest tests/test_gpu_parity.py tests/test_gpu_dense.py -m gpu -x -q 2>&1 | tail -3
python bench.py --no-cpu-baseline --steps 8 --e2e-steps 4 > gpurun_out/q.json 2> gpurun_out/q.err || tail -5 gpurun_out/q.err
python - <<'P'
import json
d = json.load(open("gpurun_out/q.json"))
print("fps", d["value"], "e2e", d["e2e"]["value"], " ".join("%s=%.4f" % (k["stage"], k["ms_per_batch"]) for k in d["kernels"]), "knn=%.4f" % d["roofline_popc"]["ms_per_batch"])
P
