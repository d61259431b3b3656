#!/bin/bash
# GPU box: warp-per-rollout (legged_coop) vs thread-per-rollout (legged_sparse) on the quadruped at the per-GPU batch sizes
# that sharding the headline workload produces (bench nominal, 1 s horizon)
out=gpurun_out/crossover_legged.txt; : > $out
for n in 1024 2048 4096 8192 16384 32768; do
  for eng in warp thread; do
    PS_ENGINE=$eng python bench.py --task "Quadruped Flat" --samples $n --horizon 101 --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
j = json.loads(sys.stdin.read().strip().splitlines()[-1])
t = j['tasks'][0]
print('N $n engine $eng', t['kernel'], 'rollout_ms %.1f' % t['kernel_ms']['rollout'], 'sample-steps/s %.4g' % j['value'])" >> $out
  done
done
cat $out
