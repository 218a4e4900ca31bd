#!/bin/bash
# every instantiated shape mask for the closest-hit and the shadow kernel on scenes whose only analytic kind is Rect
#   bash scripts/gpu_shape_masks.sh <tag> "<workloads>"
TAG=$1; WLS=$2
for wl in $WLS; do
  for cfg in "-1 -1" "32 -1" "51 -1" "124 -1" "127 -1" "895 -1" "-1 32" "-1 51" "-1 124" "-1 127" "-1 895"; do
    set -- $cfg
    PB2_TRACE_SHAPES=$1 PB2_SHADOW_SHAPES=$2 python bench.py --steps 2 --warmup 3 --workload $wl --no-cpu-baseline --no-scenes > gpurun_out/${TAG}_$wl.json 2> gpurun_out/${TAG}_$wl.err
    python - <<PY
import json
d=json.loads(open("gpurun_out/${TAG}_$wl.json").readline()); r=d["roofline"]; s=d["kernel_time_shares"]
print("$wl trace_mask $1 shadow_mask $2: Msamples/s %.1f  in-kernel closest %.0f  shares trace %.3f shadow %.3f  film_sum %s" % (d["value"], r["mrays_per_s_in_kernel"], s["trace"], s["shadow"], d["details"]["film_sum_rgb"]))
PY
  done
done
