#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O; T=r2s18
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 3 python scripts/sanitize_small.py > $O/${T}_memcheck.log 2>&1; echo "memcheck rc=$?"; tail -4 $O/${T}_memcheck.log
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 3 python scripts/sanitize_small.py > $O/${T}_racecheck.log 2>&1; echo "racecheck rc=$?"; tail -4 $O/${T}_racecheck.log
bash scripts/gpu_variants.sh $T "c1 c5" main 2>&1 | tee $O/${T}_variants.txt
