#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O; T=r2s5
cp profiles/b_ray.json $O/b_ray.json
( time PB2_B_RAY_OUT=$O/b_ray.json timeout 1200 python scripts/make_b_ray.py --recorded c1 c3 c4 c4_field c5 ) > $O/${T}_b_ray.log 2>&1; echo "b_ray rc=$?"; tail -4 $O/${T}_b_ray.log
cp $O/b_ray.json profiles/b_ray.json
( time timeout 900 python bench.py --steps 5 --warmup 3 ) > $O/${T}_bench_c5.json 2> $O/${T}_bench_c5.err; echo "bench rc=$?"; tail -3 $O/${T}_bench_c5.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2s5_bench_c5.json').readline())
print(d['value'], d['e2e']['value'], d['e2e_cold'], d['roofline']['frac'], d.get('measured_read_gbs'))
for k,v in d.get('scenes',{}).items(): print(k, v['value'], v['e2e'], v['e2e_cold_s'], v['mrays_per_s'], v['rays_per_sample'], v['roofline'].get('frac'), v['roofline'].get('bound'), v.get('cpu_baseline',{}).get('value'), v['closest_fallback_share'])
print(d.get('cpu_baseline'))
PY
( time timeout 600 python bench.py --impl reference --steps 2 --warmup 1 ) > $O/${T}_bench_ref_c5.json 2> $O/${T}_bench_ref_c5.err; echo "ref rc=$?"; cat $O/${T}_bench_ref_c5.json | cut -c1-600
