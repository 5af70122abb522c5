# device frame time against ATRAY_SPLIT_RATIO16 (half-tile mode when 16 x busy tiles <= ratio x warps) per world size
for w in 1 2 4 8; do
for r in 0 16 24 32 48 64 1000000; do
  echo "world $w ATRAY_SPLIT_RATIO16=$r"; ATRAY_SPLIT_RATIO16=$r python tools/frame_probe.py --reps 100 --world $w | tail -1
done
done
