#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O; T=r2s17
CMD="python bench.py --steps 1 --warmup 1 --workload c1 --no-cpu-baseline --no-scenes"
PB2_LANES=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_launches_c1.csv $CMD > $O/${T}_ncu_launches_c1.log 2>&1; echo "launches rc=$?"
python scripts/launch_shares.py $O/${T}_launches_c1.csv | tee $O/${T}_launch_shares_c1.txt | tail -16
PB2_LANES=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_trace_fallback$|k_trace$' -s 8 -c 4 -f -o $O/${T}_prof_c1 $CMD > $O/${T}_ncu_full_c1.log 2>&1; echo "full rc=$?"
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q -m gpu 2>&1 | tail -2
