#!/bin/bash
# One GPU session: parity tests, the bench line, the reference arm, and the ncu evidence.
#   gpurun --timeout 1500 -- 'bash scripts/gpu_round.sh <tag> [steps]'
# Everything lands in gpurun_out/<tag>_*; copy what should be judged into profiles/.
set -u
TAG=${1:-r1}
STEPS=${2:-tests,bench,ref,ncu}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${TAG}_smi.csv 2>&1
if [[ $STEPS == *tests* ]]; then
  python -m pytest tests -x -q -m gpu > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest_gpu.log
  tail -5 $O/${TAG}_pytest_gpu.log
fi
if [[ $STEPS == *smoke* ]]; then
  python -c 'import __graft_entry__ as g; g.smoke()' > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/${TAG}_smoke.log
fi
if [[ $STEPS == *bench* ]]; then
  python bench.py --steps 3 --warmup 3 > $O/${TAG}_bench_c5.json 2> $O/${TAG}_bench_c5.err; echo "bench rc=$?"
  cat $O/${TAG}_bench_c5.json
  for wl in c1 c3 c4; do
    python bench.py --steps 3 --warmup 3 --workload $wl --cpu-seconds 6 > $O/${TAG}_bench_$wl.json 2> $O/${TAG}_bench_$wl.err
    echo "bench $wl rc=$?"
  done
fi
if [[ $STEPS == *ref* ]]; then
  python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_ref_c5.json 2> $O/${TAG}_bench_ref_c5.err; echo "ref rc=$?"
  cat $O/${TAG}_bench_ref_c5.json
fi
if [[ $STEPS == *ncu* ]]; then
  CMD="python bench.py --steps 1 --warmup 1 --workload c5s --no-cpu-baseline"
  $CMD > $O/${TAG}_plain_c5s.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches_c5s.csv \
      $CMD > $O/${TAG}_ncu_launches.log 2>&1
  echo "ncu launches rc=$?"
  python scripts/launch_shares.py $O/${TAG}_launches_c5s.csv > $O/${TAG}_launch_shares_c5s.txt
  # the default bench command (C5): one whole render = 32 waves x ~51 launches
  CMD5="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
  ncu --metrics gpu__time_duration.sum --clock-control none -c 1700 --csv --log-file $O/${TAG}_launches_c5.csv \
      $CMD5 > $O/${TAG}_ncu_launches_c5.log 2>&1
  echo "ncu launches c5 rc=$?"
  python scripts/launch_shares.py $O/${TAG}_launches_c5.csv > $O/${TAG}_launch_shares_c5.txt; tail -2 $O/${TAG}_launch_shares_c5.txt
  $CMD > $O/${TAG}_plain_c5s.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:'k_trace$|k_shadow$|k_shade_mat|k_classify|k_shade_terminal' -s 14 -c 7 \
      -f -o $O/${TAG}_prof_c5s $CMD > $O/${TAG}_ncu_full.log 2>&1
  echo "ncu full rc=$?"
fi
