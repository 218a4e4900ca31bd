#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O; T=r2s11
timeout 1200 python -m pytest tests -x -q -m gpu > $O/${T}_pytest_gpu.log 2>&1; echo "pytest main rc=$?"; tail -3 $O/${T}_pytest_gpu.log
PINE_B200_LIB=$PWD/pine_b200/lib_uf/libpine_b200.so timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_scale_parity.py -x -q -m gpu > $O/${T}_pytest_uf.log 2>&1; echo "pytest uf rc=$?"; tail -3 $O/${T}_pytest_uf.log
bash scripts/gpu_variants.sh $T "c5 c4 c1 c3" main uf 2>&1 | tee $O/${T}_variants.txt
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r2s11_main_c5.json").readline()); print("cold", d["e2e_cold"]["seconds"], d["details"]["host_bvh_build_ms"], d["details"]["upload_scene_s"], "e2e", d["e2e"]["value"])
PY
