#!/bin/bash
# round 2, session 1: new parity tests + lanes-in-flight A/B
set -u
O=gpurun_out; mkdir -p $O; T=r2s1
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${T}_smi.csv 2>&1
nproc > $O/${T}_nproc.txt
timeout 900 python -m pytest tests -x -q -m gpu > $O/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/${T}_pytest_gpu.log
tail -5 $O/${T}_pytest_gpu.log
for wl in c5 c1 c3; do
  for lanes in 1 2; do
    PB2_LANES=$lanes timeout 600 python bench.py --steps 3 --warmup 3 --workload $wl --no-cpu-baseline > $O/${T}_bench_${wl}_l$lanes.json 2> $O/${T}_bench_${wl}_l$lanes.err
    echo "$wl lanes=$lanes rc=$? $(python -c "import json;d=json.load(open('$O/${T}_bench_${wl}_l$lanes.json'));print(d['value'],d['e2e']['value'],d['kernel_time_shares'])")"
  done
done
for tb in 6 7; do
  PB2_LANES=2 PB2_TRACE_BLOCKS=$tb timeout 600 python bench.py --steps 3 --warmup 3 --workload c5 --no-cpu-baseline > $O/${T}_bench_c5_l2_tb$tb.json 2> $O/${T}_bench_c5_l2_tb$tb.err
  echo "c5 lanes=2 tb=$tb rc=$? $(python -c "import json;d=json.load(open('$O/${T}_bench_c5_l2_tb$tb.json'));print(d['value'],d['e2e']['value'])")"
done
