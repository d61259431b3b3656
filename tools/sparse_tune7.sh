#!/bin/bash
# GPU box: lockstep granularity of the quadruped thread-per-rollout kernel (runtime switches, bits 16-18 of PS_COOP_SERIAL)
out=gpurun_out/sparse_tune7.txt; : > $out
for v in 0 65536 131072 196608 262144 327680; do
  PS_COOP_SERIAL=$v python tools/profile_kernel.py --task "Quadruped Flat" --N 65536 --H 61 --launches 2 --nominal random | grep "launch 1" | sed "s/^/switch $v H61 /" | cut -c1-140 >> $out
  PS_COOP_SERIAL=$v python tools/profile_kernel.py --task "Quadruped Flat" --N 65536 --H 11 --launches 2 | grep "launch 1" | sed "s/^/switch $v H11 /" | cut -c1-140 >> $out
done
cat $out
