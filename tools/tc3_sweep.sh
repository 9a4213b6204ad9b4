#!/bin/bash
# Development sweep (GPU box): v2 kernel vs the v3 configurations of mlp_tc_host.inl::launch3_cfg.
# usage: tools/tc3_sweep.sh [cfg list]   -> gpurun_out/tc3_sweep.log
mkdir -p gpurun_out
: > gpurun_out/tc3_sweep.log
echo "== v2" >> gpurun_out/tc3_sweep.log
NSDF_TC_KERNEL=2 timeout 200 python tools/tc_probe.py --n 4194304 --precs fp16,fp16x3 >> gpurun_out/tc3_sweep.log 2>&1
for c in ${@:-0 1 2 3 4 5 6 7}; do
  echo "== v3 cfg $c" >> gpurun_out/tc3_sweep.log
  NSDF_TC_KERNEL=3 NSDF_TC3_CFG=$c timeout 200 python tools/tc_probe.py --n 4194304 --precs fp16,bf16,fp16x3 >> gpurun_out/tc3_sweep.log 2>&1 || echo "FAILED rc=$?" >> gpurun_out/tc3_sweep.log
done
grep -E "^==|lattice_gqps|FAILED|rror" gpurun_out/tc3_sweep.log | sed -E 's/"err_p99.*"lattice_ms_best"/... "lattice_ms_best"/' | cut -c1-260
