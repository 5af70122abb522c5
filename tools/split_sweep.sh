# device frame time against the half-tile knobs, per world size and rank:
# ATRAY_SPLIT_RATIO16 (half-tile mode when 16 x busy tiles <= ratio x warps; 0 = never) and
# ATRAY_SPLIT_MIN_CLASS (in that mode only tiles of this cost class and above are halved)
for wr in "8 0" "8 3" "8 5" "4 0" "4 2" "2 0" "1 0"; do
set -- $wr
for mc in 8 7 6 5 4 2 0; do
  echo -n "world $1 rank $2 always-split min_class=$mc: "; ATRAY_SPLIT_RATIO16=100000000 ATRAY_SPLIT_MIN_CLASS=$mc python tools/frame_probe.py --reps 50 --world $1 --rank $2 | tail -1
done
done
