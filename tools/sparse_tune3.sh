#!/bin/bash
# GPU box: code-size variants of the thread-per-rollout quadruped kernel (instruction-delivery bound)
out=gpurun_out/sparse_tune3.txt; : > $out
for cfg in "0" "1" "2"; do
  touch mujoco_mpc_b200/csrc/ps_variant_legged_sparse.cu
  python -c "from mujoco_mpc_b200 import build; build.build(extra=['-DPS_COMPACT_LOOPS=$cfg'])" > /dev/null 2>&1 || { echo "build failed $cfg" >> $out; continue; }
  for per in 0 2; do
    r=$(PS_SPARSE_BLOCKS_PER_SM=$per python tools/profile_kernel.py --task "Quadruped Flat" --N 65536 --H 31 --launches 2 --nominal random 2>&1 | grep "launch 1" | sed 's/.*kernel_ms=\([0-9.]*\).*sample-steps.s=\([0-9.e+]*\).*/\1 ms \2/')
    echo "compact $cfg blocks_per_sm $per : $r" >> $out
  done
done
cat $out
