#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O; T=r2s6
timeout 1200 python -m pytest tests -x -q -m gpu > $O/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${T}_pytest_gpu.log
bash scripts/gpu_variants.sh $T "c5 c4 c1" base main smem73 smem585 2>&1 | tee $O/${T}_variants.txt
