#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O; T=r2s8
timeout 1200 python -m pytest tests -x -q -m gpu > $O/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 $O/${T}_pytest_gpu.log
bash scripts/gpu_variants.sh $T "c1 c3" main smem0 main smem0 2>&1 | tee $O/${T}_variants.txt
