#!/bin/bash
# GPU box: headline quadruped workload (65,536 x 1 s, bench nominal) under lockstep / block variants
out=gpurun_out/sparse_tune5.txt; : > $out
for cfg in "1 512" "0 128"; do
  set -- $cfg
  touch mujoco_mpc_b200/csrc/ps_variant_legged_sparse.cu
  python -c "from mujoco_mpc_b200 import build; build.build(extra=['-DPS_SPARSE_LOCKSTEP=$1','-DPS_SPARSE_BLOCK=$2'])" > /dev/null 2>&1 || { echo "build failed $cfg" >> $out; continue; }
  python bench.py --workload quadruped_65536_1s --steps 2 --warmup 1 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
j = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('lockstep $1 block $2', '%.4g' % j['value'], [(t['kernel'], round(t['kernel_ms']['rollout'], 1)) for t in j['tasks']], 'e2e device_ms', round(j['e2e'].get('device_ms', 0), 1))" >> $out
done
cat $out
