#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O; T=r2s16
timeout 1200 python -m pytest tests -x -q -m gpu > $O/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${T}_pytest_gpu.log
PB2_VERBOSE=1 timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-scenes > $O/${T}_bench_c5.json 2> $O/${T}_bench_c5.err; echo "bench rc=$?"
grep -E "pb2_upload_scene|build_wide" $O/${T}_bench_c5.err | head -12
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r2s16_bench_c5.json").readline()); print("value", d["value"], "cold", d["e2e_cold"]["seconds"], "build", d["details"]["host_bvh_build_ms"], "upload", d["details"]["upload_scene_s"], "e2e", d["e2e"]["value"])
PY
