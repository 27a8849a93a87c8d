#!/bin/bash
# Round artefacts on the GPU box: gpu_profile.sh (tests, both bench arms, launch list, splat capture)
# + a full capture of the assembly kernel + the secondary workload lines.  TAG=$1
TAG=${1:-r1}
OUT=gpurun_out
bash scripts/gpu_profile.sh $TAG > $OUT/${TAG}_profile.log 2>&1
SHORT="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
ncu --set full --clock-control none --import-source on -k regex:assemble_kernel -s 3 -c 1 \
    -f -o $OUT/${TAG}_asm $SHORT > $OUT/${TAG}_ncu_asm.log 2>&1
for w in crowd_1281_50p dense_641_5p randinit_641_tamed randinit_641; do
  timeout 900 python bench.py --workload $w --steps 5 > $OUT/${TAG}_$w.json 2> $OUT/${TAG}_$w.err
done
timeout 600 python bench.py --workload cifdet_641_20b --steps 10 > $OUT/${TAG}_cifdet.json 2> $OUT/${TAG}_cifdet.err
tail -2 $OUT/${TAG}_pytest_gpu.log; tail -c 600 $OUT/${TAG}_bench.json
