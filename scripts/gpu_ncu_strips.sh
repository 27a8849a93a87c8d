#!/bin/bash
# ncu --set full of the strips kernel (default and crowd workloads).  TAG=$1, extra bench args in $2
TAG=${1:-r2x}; OUT=gpurun_out; mkdir -p $OUT
for wl in coco17_641_5p crowd_1281_50p; do
  CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --value-only --distinct 1 --workload $wl ${2:-}"
  $CMD > $OUT/${TAG}_plain_$wl.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:'cifhr_strips' -s 2 -c 1 \
      -f -o $OUT/${TAG}_$wl $CMD > $OUT/${TAG}_ncu_$wl.log 2>&1
  tail -2 $OUT/${TAG}_ncu_$wl.log
done
