#!/bin/bash
# GPU box: lockstep block size of the quadruped thread-per-rollout kernel on the final tree (runtime switch, no rebuild)
out=gpurun_out/sparse_tune8.txt; : > $out
run() {
  env $1 python bench.py --workload quadruped_65536_1s --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
j = json.loads(sys.stdin.read().strip().splitlines()[-1]); t = j['tasks'][0]
print('[$1]', t['kernel'], 'rollout_ms %.1f' % t['kernel_ms']['rollout'], 'value %.4g' % j['value'])" >> $out
}
run "PS_X=0"
run "PS_SPARSE_BLOCK_RT=224"
run "PS_SPARSE_BLOCK_RT=320"
run "PS_SPARSE_BLOCK_RT=128"
cat $out
