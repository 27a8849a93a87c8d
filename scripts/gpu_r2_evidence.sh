#!/bin/bash
# ncu evidence for round 2 (one GPU): launch list of the default step + one `--set full` capture of EVERY
# kernel of a decode on the default, crowd, dense and tamed random-init workloads + PCIe bytes of the two
# kernels that read pinned host memory on the host-mapped route.  TAG=$1
set -u
TAG=${1:-r2n}; OUT=gpurun_out; mkdir -p $OUT
K='fields_scan|cifhr_strips|cifhr_splat|seeds_field_sort|seeds_merge|seeds_from|seeds_collect|seeds_sort|caf_score_bin|caf_scored|caf_bin|assemble_kernel|pack_results|nms_kernel'
BASE="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --value-only --distinct 1"
$BASE > $OUT/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv \
    --log-file $OUT/${TAG}_launches.csv $BASE > $OUT/${TAG}_ncu_launches.log 2>&1
for wl in coco17_641_5p crowd_1281_50p dense_641_5p randinit_641_tamed; do
  CMD="$BASE --workload $wl"
  $CMD > $OUT/${TAG}_plain_$wl.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:"$K" -s 16 -c 8 \
      -f -o $OUT/${TAG}_$wl $CMD > $OUT/${TAG}_ncu_$wl.log 2>&1
  tail -2 $OUT/${TAG}_ncu_$wl.log
done
HOST="python scripts/stream_sweep.py 32x4"
$HOST > $OUT/${TAG}_plain_host.log 2>&1 &&
ncu --metrics pcie__read_bytes.sum,pcie__write_bytes.sum,gpu__time_duration.sum --clock-control none \
    -k regex:'fields_scan|caf_score_bin' -s 40 -c 16 --csv --log-file $OUT/${TAG}_pcie.csv $HOST > $OUT/${TAG}_ncu_pcie.log 2>&1
tail -3 $OUT/${TAG}_ncu_pcie.log
ls -la $OUT | grep $TAG
