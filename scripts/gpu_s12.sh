#!/bin/bash
set -u
N=${1:-2}
O=gpurun_out; mkdir -p $O; T=r2s12_n$N
NCCL_DEBUG=INFO timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > $O/${T}_bench_c5.json 2> $O/${T}_bench_c5.err; echo "torchrun bench rc=$?"
wc -l $O/${T}_bench_c5.json; grep -c "NCCL INFO" $O/${T}_bench_c5.err; grep -m3 "nranks\|Init COMPLETE" $O/${T}_bench_c5.err | cut -c1-200
python - <<PY
import json
d=json.loads(open("$O/${T}_bench_c5.json").readline())
print("N=$N value", d["value"], "e2e", d["e2e"]["value"], "cold", d["e2e_cold"]["seconds"], "h2d", d["e2e"]["h2d_bytes_per_step"], "build_ms", d["details"]["host_bvh_build_ms"])
PY
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 1 --warmup 0 --ref-step-seconds 2 > $O/${T}_bench_ref.json 2> $O/${T}_bench_ref.err; echo "ref arm rc=$? lines=$(wc -l < $O/${T}_bench_ref.json)"
for mode in samples rows; do timeout 900 python scripts/bench_multi.py --workload c5 --steps 3 --mode $mode > $O/${T}_multi_$mode.json 2> $O/${T}_multi_$mode.err; echo "multi $mode rc=$?"; cat $O/${T}_multi_$mode.json; done
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q -m gpu 2>&1 | tail -2
