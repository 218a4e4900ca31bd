#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O; T=${1:-r2s3}
timeout 1200 python -m pytest tests -x -q -m gpu > $O/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 $O/${T}_pytest_gpu.log
