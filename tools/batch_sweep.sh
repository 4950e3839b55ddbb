for b in 16 32 64 128; do
python bench.py --batch $b --no-cpu-baseline --steps 8 --e2e-steps 2 --e2e-seconds 0.1 --sustain 0 --cfg3-frames 512 --cfg5-frames 2 > gpurun_out/qb.json 2> gpurun_out/qb.err || tail -3 gpurun_out/qb.err
python - $b <<'P'
import json,sys
d=json.load(open("gpurun_out/qb.json")); B=int(sys.argv[1])
print("B", B, "fps", d["value"], " ".join("%s=%.3f"%(k["stage"],k["us_per_frame"]) for k in d["kernels"]), "knn=%.3f"%d["roofline_popc"]["us_per_frame"])
P
done
