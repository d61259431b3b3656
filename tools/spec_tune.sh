#!/bin/bash
# GPU box: rebuild the model-specialised tiny kernels under a few (block, resident blocks) settings and time the
# cartpole 65,536 x 1 s and acrobot+particle 4,096 workloads.  Output: gpurun_out/spec_tune.txt
out=gpurun_out/spec_tune.txt; : > $out
for cfg in "128 4" "128 3" "128 5" "64 8" "256 2" "64 6"; do
  set -- $cfg
  touch mujoco_mpc_b200/csrc/ps_variant_tiny_spec.cu
  python -c "from mujoco_mpc_b200 import build; build.build(extra=['-DPS_SPEC_BLOCK=$1','-DPS_SPEC_MINBLOCKS=$2'])" > /dev/null 2>&1 || { echo "build failed $cfg" >> $out; continue; }
  for wl in cartpole_65536_1s acrobot+particle_4096_cubic; do
    python bench.py --workload $wl --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
j = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('block $1 minblocks $2', '$wl', '%.4g' % j['value'], [(t['kernel'], round(t['kernel_ms']['rollout'], 4)) for t in j['tasks']])" >> $out
  done
done
# restore the default build
touch mujoco_mpc_b200/csrc/ps_variant_tiny_spec.cu
python -c "from mujoco_mpc_b200 import build; build.build()" > /dev/null 2>&1
cat $out
