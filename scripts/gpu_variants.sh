#!/bin/bash
# bench.py across build variants:  bash scripts/gpu_variants.sh <tag> "<workloads>" <variant suffixes...>
# ("" = the product library). Variants are built with `make -C pine_b200/csrc VARIANT=_x EXTRA=-D...`.
TAG=$1; WLS=$2; shift 2
for v in "$@"; do
  if [ "$v" = "main" ]; then unset PINE_B200_LIB; else export PINE_B200_LIB=$PWD/pine_b200/lib_$v/libpine_b200.so; fi
  for wl in $WLS; do
    python bench.py --steps 3 --warmup 3 --workload $wl --no-cpu-baseline --no-scenes > gpurun_out/${TAG}_${v}_$wl.json 2> gpurun_out/${TAG}_${v}_$wl.err
    python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${TAG}_${v}_$wl.json").readline())
    s=d["kernel_time_shares"]; r=d["roofline"]
    print("$v", "$wl", "Msamples/s %.1f Mrays/s %.1f ms %.2f | in-kernel closest Mrays/s %.0f frac %.3f | shares trace %.2f shadow %.2f shade %.2f" % (d["value"], d["mrays_per_s"], d["ms_per_step"], r["mrays_per_s_in_kernel"], r["frac"] or 0, s["trace"], s["shadow"], s["shade"]), "film_sum", d["details"]["film_sum_rgb"])
except Exception as e:
    print("$v $wl failed", e); print(open("gpurun_out/${TAG}_${v}_$wl.err").read()[-800:])
PY
  done
done
