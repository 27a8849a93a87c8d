#!/bin/bash
# ncu evidence for round 2 (one GPU): launch list of the default step + one `--set full` capture of EVERY
# kernel of a decode on the default, crowd and dense workloads, summarised ON THE BOX (the reports are too
# large to travel: only the default workload's comes back) + PCIe bytes of the kernels that read pinned host
# memory on the host-mapped route.  TAG=$1
set -u
TAG=${1:-r2n}; OUT=gpurun_out; TMP=/tmp/ncu_$TAG; mkdir -p $OUT $TMP
K='fields_scan|cifhr_strips|cifhr_splat|seeds_field_sort|seeds_merge|seeds_from|seeds_collect|seeds_sort|caf_score_bin|caf_scored|caf_bin|assemble_kernel|pack_results|nms_kernel'
BASE="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --value-only --distinct 1"
$BASE > $OUT/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv \
    --log-file $OUT/${TAG}_launches.csv $BASE > $OUT/${TAG}_ncu_launches.log 2>&1
ARGS=""
for wl in coco17_641_5p crowd_1281_50p dense_641_5p ${EXTRA_WL:-}; do
  CMD="$BASE --workload $wl"
  $CMD > $OUT/${TAG}_plain_$wl.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:"$K" -s 16 -c 8 \
      -f -o $TMP/${TAG}_$wl $CMD > $OUT/${TAG}_ncu_$wl.log 2>&1
  tail -1 $OUT/${TAG}_ncu_$wl.log
  ARGS="$ARGS $wl:$TMP/${TAG}_$wl.ncu-rep"
  ncu -i $TMP/${TAG}_$wl.ncu-rep --page raw --csv > $OUT/${TAG}_raw_$wl.csv 2>/dev/null
done
python scripts/summarize_ncu_multi.py $OUT/${TAG}_kernels.md $ARGS > $OUT/${TAG}_summarize.log 2>&1
for k in cifhr_strips assemble_kernel fields_scan; do
  ncu -i $TMP/${TAG}_coco17_641_5p.ncu-rep -k regex:$k --page source --csv --print-source cuda > $OUT/${TAG}_source_$k.csv 2>/dev/null
done
gzip -f $OUT/${TAG}_source_*.csv $OUT/${TAG}_raw_*.csv
cp $TMP/${TAG}_coco17_641_5p.ncu-rep $OUT/ 2>/dev/null
HOST="python scripts/stream_sweep.py 32x4"
ncu --metrics pcie__read_bytes.sum,pcie__write_bytes.sum,gpu__time_duration.sum --clock-control none \
    -k regex:'fields_scan|caf_score_bin' -s 40 -c 16 --csv --log-file $OUT/${TAG}_pcie.csv $HOST > $OUT/${TAG}_ncu_pcie.log 2>&1
tail -3 $OUT/${TAG}_pcie.csv | cut -c1-300
du -sh $OUT
