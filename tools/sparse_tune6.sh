#!/bin/bash
# GPU box: headline quadruped workload (65,536 x 1 s, bench nominal) under runtime block size / residency caps (no rebuild)
out=gpurun_out/sparse_tune6.txt; : > $out
run() {
  env $1 python bench.py --workload quadruped_65536_1s --steps 2 --warmup 1 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
j = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$1', '%.4g' % j['value'], [(t['kernel'], round(t['kernel_ms']['rollout'], 1)) for t in j['tasks']])" >> $out
}
run "PS_X=0"
run "PS_SPARSE_BLOCK_RT=256"
run "PS_SPARSE_BLOCK_RT=256 PS_SPARSE_BLOCKS_PER_SM=1"
run "PS_SPARSE_BLOCK_RT=128 PS_SPARSE_BLOCKS_PER_SM=1"
run "PS_SPARSE_BLOCK_RT=128 PS_SPARSE_BLOCKS_PER_SM=2"
run "PS_ENGINE=warp"
cat $out
