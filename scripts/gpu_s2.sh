#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O; T=r2s2
timeout 600 python -m pytest tests/test_gpu_scale_parity.py -x -q -m gpu -k "wave_plan or three_sigma" > $O/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${T}_pytest.log
timeout 600 python scripts/explore_shards.py > $O/${T}_shards.txt 2>&1; cat $O/${T}_shards.txt
run() { # name env...
  local name=$1; shift
  env "$@" timeout 600 python bench.py --steps 3 --warmup 3 --workload $WL --no-cpu-baseline > $O/${T}_bench_${WL}_$name.json 2> $O/${T}_bench_${WL}_$name.err
  echo "$WL $name rc=$? $(python -c "import json;d=json.load(open('$O/${T}_bench_${WL}_$name.json'));print(round(d['value'],1),round(d['e2e']['value'],1))")"
}
WL=c5; run l2 PB2_LANES=2; run l3 PB2_LANES=3; run l4 PB2_LANES=4; run l1a PB2_LANES=1 PB2_SHADOW_ASYNC=1; run l2a PB2_LANES=2 PB2_SHADOW_ASYNC=1; run l3a PB2_LANES=3 PB2_SHADOW_ASYNC=1
WL=c1; run l2 PB2_LANES=2; run l4 PB2_LANES=4; run l2a PB2_LANES=2 PB2_SHADOW_ASYNC=1; run l4a PB2_LANES=4 PB2_SHADOW_ASYNC=1
WL=c3; run l2 PB2_LANES=2; run l4 PB2_LANES=4; run l2a PB2_LANES=2 PB2_SHADOW_ASYNC=1
