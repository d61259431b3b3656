#!/bin/bash
# GPU box: block-lockstep thread-per-rollout quadruped kernel vs free-running, a few block sizes
out=gpurun_out/sparse_tune4.txt; : > $out
python -m pytest tests/test_gpu_parity.py -q -x -k "quadruped or Quadruped or overflow" > gpurun_out/lockstep_tests.log 2>&1; tail -3 gpurun_out/lockstep_tests.log >> $out
for cfg in "1 128 512" "0 128 512" "1 256 512" "1 512 512" "1 64 512" "1 256 768"; do
  set -- $cfg
  touch mujoco_mpc_b200/csrc/ps_variant_legged_sparse.cu
  python -c "from mujoco_mpc_b200 import build; build.build(extra=['-DPS_SPARSE_LOCKSTEP=$1','-DPS_SPARSE_BLOCK=$2','-DPS_SPARSE_RESIDENT=$3'])" > /dev/null 2>&1 || { echo "build failed $cfg" >> $out; continue; }
  r=$(python tools/profile_kernel.py --task "Quadruped Flat" --N 65536 --H 31 --launches 2 --nominal random 2>&1 | grep "launch 1" | sed 's/.*kernel_ms=\([0-9.]*\).*sample-steps.s=\([0-9.e+]*\).*/\1 ms \2/')
  echo "lockstep $1 block $2 resident $3 : $r" >> $out
done
cat $out
