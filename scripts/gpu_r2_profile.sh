#!/bin/bash
# Round-2 evidence run on the GPU box (under gpurun): parity tests, the default bench line, then ncu
# captures of (a) every kernel of the default step and (b) the three long kernels of the crowd
# workload.  All ncu runs of one call count as one tool use.  TAG=$1
set -u
TAG=${1:-r2a}
OUT=gpurun_out
mkdir -p $OUT
python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $OUT/${TAG}_pytest_gpu.log
tail -2 $OUT/${TAG}_pytest_gpu.log
python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
echo "bench rc=$?" >> $OUT/${TAG}_bench.err
python bench.py --workload crowd_1281_50p --steps 32 --no-cpu-baseline > $OUT/${TAG}_crowd.json 2> $OUT/${TAG}_crowd.err
SHORT="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$SHORT > $OUT/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file $OUT/${TAG}_launches.csv $SHORT > $OUT/${TAG}_ncu_launches.log 2>&1
# every kernel of the default step, one launch each (warm-up launches skipped per kernel with -s)
$SHORT > $OUT/${TAG}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on \
    -k regex:'cif_compact|caf_scored|caf_bin|seeds_from|seeds_sort|nms_kernel|assemble_kernel|cifhr_splat' \
    -s 24 -c 8 -f -o $OUT/${TAG}_step $SHORT > $OUT/${TAG}_ncu_step.log 2>&1
CROWD="python bench.py --workload crowd_1281_50p --steps 2 --warmup 3 --no-cpu-baseline"
$CROWD > $OUT/${TAG}_plain3.log 2>&1 &&
ncu --set full --clock-control none --import-source on \
    -k regex:'seeds_sort|assemble_kernel|cifhr_splat|cif_compact' \
    -s 12 -c 4 -f -o $OUT/${TAG}_crowd $CROWD > $OUT/${TAG}_ncu_crowd.log 2>&1
tail -3 $OUT/${TAG}_ncu_step.log $OUT/${TAG}_ncu_crowd.log
tail -c 400 $OUT/${TAG}_bench.json
