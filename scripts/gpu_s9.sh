#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O; T=r2s9
timeout 1200 python -m pytest tests -x -q -m gpu > $O/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 $O/${T}_pytest_gpu.log
bash scripts/gpu_variants.sh $T "c1 c5" main 2>&1 | tee $O/${T}_variants.txt
python - <<'PY'
import json
for wl in ("c1","c5"):
    d=json.loads(open(f"gpurun_out/r2s9_main_{wl}.json").readline()); print(wl, "fallback share", d["closest_fallback_share"])
PY
