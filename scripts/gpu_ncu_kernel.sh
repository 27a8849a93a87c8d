#!/bin/bash
# bench + one ncu --set full capture of the kernel matching $2 (regex); TAG=$1
TAG=${1:-k}; PAT=${2:-assemble_kernel}
OUT=gpurun_out
python bench.py --no-cpu-baseline > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; tail -c 1500 $OUT/${TAG}_bench.json
SHORT="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$SHORT > $OUT/${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$PAT -s 3 -c 1 \
    -f -o $OUT/${TAG}_cap $SHORT > $OUT/${TAG}_ncu.log 2>&1
tail -3 $OUT/${TAG}_ncu.log
