#!/bin/bash
# Env-knob sweep of bench.py on one workload:  bash scripts/gpu_sweep.sh <tag> <workload> VAR v1 v2 ...
# Writes gpurun_out/<tag>_<VAR>_<v>.json and prints a one-line summary per value.
TAG=$1; WL=$2; VAR=$3; shift 3
mkdir -p gpurun_out
for v in "$@"; do
  env $VAR=$v python bench.py --steps 3 --warmup 3 --workload $WL --no-cpu-baseline > gpurun_out/${TAG}_${VAR}_$v.json 2> gpurun_out/${TAG}_${VAR}_$v.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/${TAG}_${VAR}_$v.json"))
    s=d["kernel_time_shares"]; r=d["roofline"]
    print("$VAR=$v", "$WL", "Msamples/s %.1f Mrays/s %.1f ms %.2f | in-kernel closest Mrays/s %.0f frac %.3f | shares trace %.2f shadow %.2f shade %.2f" % (d["value"], d["mrays_per_s"], d["ms_per_step"], r["mrays_per_s_in_kernel"], r["frac"] or 0, s["trace"], s["shadow"], s["shade"]), "film_sum", d["config"]["film_sum_rgb"])
except Exception as e:
    print("$VAR=$v failed", e); print(open("gpurun_out/${TAG}_${VAR}_$v.err").read()[-800:])
PY
done
