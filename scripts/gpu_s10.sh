#!/bin/bash
# full bench line + reference arms + ncu evidence on the C5 size (round 2)
set -u
O=gpurun_out; mkdir -p $O; T=${1:-r2s10}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${T}_smi.csv 2>&1
( time timeout 900 python bench.py --steps 5 --warmup 3 ) > $O/${T}_bench_c5.json 2> $O/${T}_bench_c5.err; echo "bench rc=$?"; tail -3 $O/${T}_bench_c5.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/${T}_bench_ref_c5.json 2> $O/${T}_bench_ref_c5.err; echo "ref rc=$?"
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-scenes"
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 3400 --csv --log-file $O/${T}_launches_c5.csv $CMD > $O/${T}_ncu_launches_c5.log 2>&1
echo "ncu launches rc=$?"
python scripts/launch_shares.py $O/${T}_launches_c5.csv > $O/${T}_launch_shares_c5.txt; tail -8 $O/${T}_launch_shares_c5.txt
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'k_trace$|k_shadow$' -s 22 -c 11 -f -o $O/${T}_prof_c5 $CMD > $O/${T}_ncu_full_c5.log 2>&1
echo "ncu full rc=$?"; ls -la $O/${T}_prof_c5.ncu-rep
