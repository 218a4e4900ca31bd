#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O; T=r2s7
for rep in a b; do
bash scripts/gpu_variants.sh ${T}$rep "c5 c4" main smem9 smem73 smem201 2>&1 | tee -a $O/${T}_variants.txt
done
