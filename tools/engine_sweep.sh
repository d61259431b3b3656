# throughput of the three kernel flavours over the batch size (diagnostics; PS_ENGINE pins a flavour)
for task in "Quadruped Flat" "Walker"; do
  for eng in warp thread dense; do
    for N in 64 1024 8192 65536; do
      H=21; [ "$task" = "Walker" ] && H=51
      echo -n "$task engine=$eng N=$N : "
      PS_ENGINE=$eng python tools/profile_kernel.py --task "$task" --N $N --H $H --launches 2 2>&1 | grep "launch 1" | sed 's/.*variant=/variant=/'
    done
  done
done
