#!/bin/bash
# bench.py at N = 1,2,4,8 on one box (run under `gpurun --gpus 8`): writes gpurun_out/<tag>_n<N>.json
TAG=${1:-scale}; WL=${2:-c5}; NS=${3:-"1 2 4 8"}
mkdir -p gpurun_out
port=29600
for n in $NS; do
  port=$((port+1))
  if [ "$n" = "1" ]; then
    python bench.py --gpus 1 --steps 3 --warmup 3 --workload $WL --no-cpu-baseline > gpurun_out/${TAG}_n$n.json 2> gpurun_out/${TAG}_n$n.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --steps 3 --warmup 3 --workload $WL > gpurun_out/${TAG}_n$n.json 2> gpurun_out/${TAG}_n$n.err
  fi
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${TAG}_n$n.json").read().strip().splitlines()[-1])
    print("N=$n", "$WL", "Msamples/s %.1f  Mrays/s %.1f  ms/step %.2f  e2e %.1f" % (d["value"], d["mrays_per_s"], d["ms_per_step"], d["e2e"]["value"]))
except Exception as e:
    print("N=$n failed", e); print(open("gpurun_out/${TAG}_n$n.err").read()[-1500:])
PY
done
