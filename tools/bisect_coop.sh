for mask in 0 511 1 2 4 8 16 32 64 128 256; do
  echo "== PS_COOP_SERIAL=$mask"
  PS_COOP_SERIAL=$mask python tools/gpu_mismatch.py --task Walker --N 1024 --repeat 2 2>&1 | grep -E "^rep|identical"
done
