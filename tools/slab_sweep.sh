# e2e frame time against slab plans (tools/frame_probe.py --host); every plan is run twice, interleaved
for rep in 1 2; do
for cs in 1 2; do
export ATRAY_COPY_STREAMS=$cs
echo "copy streams $cs"
for w in "2,6,8,16" "2,6,8,8,6,2" "2,6,8,10,4,2" "1,3,4,8,8,4,3,1" "2,6,8,12,4" "2,6,10,10,4" "2,6,8,8,8"; do
  ATRAY_SLAB_WEIGHTS=$w python tools/frame_probe.py --host --reps 100 | tail -1
done
done
done
