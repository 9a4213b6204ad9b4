#!/bin/bash
# Development sweep (GPU box): v2 / v3 / v4 schedules of the tensor-core rungs.  -> gpurun_out/tc4_sweep.log
mkdir -p gpurun_out
L=gpurun_out/tc4_sweep.log
: > $L
echo "== v2" >> $L
NSDF_TC_KERNEL=2 timeout 200 python tools/tc_probe.py --n 4194304 --precs fp16,fp16x3 >> $L 2>&1
for c in ${@:-0 1 2 3 4}; do
  echo "== v4 cfg $c" >> $L
  NSDF_TC_KERNEL=4 NSDF_TC4_CFG=$c timeout 200 python tools/tc_probe.py --n 4194304 --precs fp16,bf16,fp16x3 >> $L 2>&1 || echo "FAILED rc=$?" >> $L
done
grep -E "^==|lattice_gqps|FAILED|rror" $L | sed -E 's/\{"err_max": ([0-9.e-]+),.*"lattice_gqps": ([0-9.]+),.*"points_gqps": ([0-9.]+).*/err \1 lattice \2 Gq\/s points \3 Gq\/s/'
