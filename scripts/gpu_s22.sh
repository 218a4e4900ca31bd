#!/bin/bash
# wave size sweep on C5: PB2_MAX_SLOTS 2^23 .. 2^26 (default 2^24), lanes 4 and 2
for cfg in "16777216 4" "8388608 4" "33554432 4" "67108864 4" "33554432 2" "67108864 2" "16777216 4"; do
  set -- $cfg
  PB2_MAX_SLOTS=$1 PB2_LANES=$2 python bench.py --steps 3 --warmup 3 --workload c5 --no-cpu-baseline --no-scenes > gpurun_out/r2s22_$1_$2.json 2> gpurun_out/r2s22_$1_$2.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2s22_$1_$2.json").readline())
    print("max_slots $1 lanes $2: Msamples/s %.1f e2e %.1f ms %.2f film_sum %s" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["details"]["film_sum_rgb"]))
except Exception as e:
    print("$1 $2 failed", e); print(open("gpurun_out/r2s22_$1_$2.err").read()[-600:])
PY
done
nvidia-smi --query-gpu=memory.used --format=csv
