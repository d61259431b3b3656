#!/bin/bash
# GPU box: rebuild the thread-per-rollout quadruped kernel under (block, resident threads) settings and time 65,536 x H=31
out=gpurun_out/sparse_tune.txt; : > $out
for cfg in "128 512" "256 768" "256 1024" "128 768" "64 512" "256 512"; do
  set -- $cfg
  touch mujoco_mpc_b200/csrc/ps_variant_legged_sparse.cu
  python -c "from mujoco_mpc_b200 import build; build.build(extra=['-DPS_SPARSE_BLOCK=$1','-DPS_SPARSE_RESIDENT=$2'])" > /dev/null 2>&1 || { echo "build failed $cfg" >> $out; continue; }
  python bench.py --task "Quadruped Flat" --samples 65536 --horizon 31 --steps 2 --warmup 1 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
j = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('block $1 resident $2', '%.4g' % j['value'], [(t['kernel'], round(t['kernel_ms']['rollout'], 2)) for t in j['tasks']])" >> $out
done
touch mujoco_mpc_b200/csrc/ps_variant_legged_sparse.cu
python -c "from mujoco_mpc_b200 import build; build.build()" > /dev/null 2>&1
cat $out
