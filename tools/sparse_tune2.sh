#!/bin/bash
# GPU box: thread-per-rollout quadruped kernel under bounded occupancy (PS_SPARSE_BLOCKS_PER_SM) x (block, register cap)
out=gpurun_out/sparse_tune2.txt; : > $out
for cfg in "128 512" "128 128" "64 64" "32 32"; do
  set -- $cfg
  touch mujoco_mpc_b200/csrc/ps_variant_legged_sparse.cu
  python -c "from mujoco_mpc_b200 import build; build.build(extra=['-DPS_SPARSE_BLOCK=$1','-DPS_SPARSE_RESIDENT=$2'])" > /dev/null 2>&1 || { echo "build failed $cfg" >> $out; continue; }
  for per in 0 1 2 3 4 6 8; do
    r=$(PS_SPARSE_BLOCKS_PER_SM=$per python tools/profile_kernel.py --task "Quadruped Flat" --N 65536 --H 31 --launches 2 --nominal random 2>&1 | grep "launch 1" | sed 's/.*kernel_ms=\([0-9.]*\).*sample-steps.s=\([0-9.e+]*\).*/\1 ms \2/')
    echo "block $1 resident $2 blocks_per_sm $per : $r" >> $out
  done
done
touch mujoco_mpc_b200/csrc/ps_variant_legged_sparse.cu
python -c "from mujoco_mpc_b200 import build; build.build()" > /dev/null 2>&1
cat $out
