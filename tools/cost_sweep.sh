# device frame time against the cost-class shift of the busy-tile lists (31 = one class = list order only)
for rep in 1 2; do
for sh in 31 0 1 2; do
  echo "ATRAY_COST_SHIFT=$sh"; ATRAY_COST_SHIFT=$sh python tools/frame_probe.py --reps 100 | tail -2
done
done
