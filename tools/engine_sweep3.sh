for task in Acrobot Particle Cartpole; do
  for eng in warp thread; do
    for N in 64 1024 4096 16384 65536; do
      H=201; [ "$task" = "Particle" ] && H=51; [ "$task" = "Cartpole" ] && H=101
      echo -n "$task engine=$eng N=$N : "
      PS_ENGINE=$eng python tools/profile_kernel.py --task "$task" --N $N --H $H --launches 3 2>&1 | grep "launch 2" | sed 's/.*variant=/variant=/'
    done
  done
done
