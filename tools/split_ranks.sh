# world 8, every rank: busy tiles per cost class (tile log, half-tile mode off) and the device frame with the half-tile
# mode off / on for classes >= 5 / >= 6
for r in 0 1 2 3 4 5 6 7; do
  ATRAY_SPLIT_RATIO16=0 ATRAY_TILE_LOG=1 python tools/tile_log_probe.py --world 8 --rank $r --brief | tail -1
  for mc in 8 6 5; do
    echo -n "   min_class=$mc: "; ATRAY_SPLIT_RATIO16=100000000 ATRAY_SPLIT_MIN_CLASS=$mc python tools/frame_probe.py --reps 50 --world 8 --rank $r | tail -1
  done
done
