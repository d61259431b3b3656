#!/bin/bash
# GPU box: which kernel should serve the planar class's big batches?  Walker, bench settings, three codings.
out=gpurun_out/planar_large_tune.txt; : > $out
run() {
  env $1 python bench.py --task Walker --samples $2 --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
j = json.loads(sys.stdin.read().strip().splitlines()[-1])
t = j['tasks'][0]
print('N $2 [$1]', t['kernel'], 'rollout_ms %.2f' % t['kernel_ms']['rollout'], 'sample-steps/s %.4g' % j['value'])" >> $out
}
for n in 4096 16384 65536; do
  run "PS_X=0" $n
  run "PS_PLANAR_LARGE=sparse PS_ENGINE=thread" $n
  run "PS_ENGINE=warp" $n
done
cat $out
