#!/bin/bash
# last set on the final build: default bench line, dense at batch 256, full ncu capture of the default step
TAG=${1:-r3z}; OUT=gpurun_out; TMP=/tmp/ncu_$TAG; mkdir -p $OUT $TMP
timeout 400 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"
timeout 500 python bench.py --workload dense_641_5p_b256 --steps 16 > $OUT/${TAG}_dense_b256.json 2> $OUT/${TAG}_dense_b256.err; echo "b256 rc=$?"; tail -3 $OUT/${TAG}_dense_b256.err
K='fields_scan|cifhr_strips|cifhr_splat|seeds_field_sort|seeds_merge|caf_score_bin|assemble_kernel|pack_results'
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --value-only --distinct 1"
ncu --set full --clock-control none --import-source on -k regex:"$K" -s 16 -c 8 -f -o $TMP/${TAG}_coco17_641_5p $CMD > $OUT/${TAG}_ncu.log 2>&1
tail -1 $OUT/${TAG}_ncu.log
python scripts/summarize_ncu_multi.py $OUT/${TAG}_kernels_default.md coco17_641_5p:$TMP/${TAG}_coco17_641_5p.ncu-rep > $OUT/${TAG}_summarize.log 2>&1
cp $TMP/${TAG}_coco17_641_5p.ncu-rep $OUT/ 2>/dev/null
du -sh $OUT
