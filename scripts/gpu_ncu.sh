#!/bin/bash
# ncu evidence for one workload: launch list + one --set full capture of the named kernels.
#   bash scripts/gpu_ncu.sh <tag> <workload> <kernel regex> <skip> <count>
TAG=$1; WL=$2; RX=$3; SKIP=${4:-6}; CNT=${5:-6}
O=gpurun_out; mkdir -p $O
CMD="python bench.py --steps 1 --warmup 1 --workload $WL --no-cpu-baseline"
$CMD > $O/${TAG}_plain_$WL.log 2>&1 || { echo "plain run failed"; tail -5 $O/${TAG}_plain_$WL.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${TAG}_launches_$WL.csv $CMD > $O/${TAG}_ncu_launches_$WL.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"$RX" -s $SKIP -c $CNT -f -o $O/${TAG}_prof_$WL $CMD > $O/${TAG}_ncu_full_$WL.log 2>&1
echo "full rc=$?"
