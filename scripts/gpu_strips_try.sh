#!/bin/bash
# strips kernel: parity, then value under the schedule knobs (cifhr_warps=-1 = the round-1 kernel)
TAG=${1:-r2u}; OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 $OUT/${TAG}_pytest.log
bash scripts/gpu_tune_sweep.sh ${TAG}_default coco17_641_5p default cifhr_warps=2 cifhr_warps=3 cifhr_warps=3,cifhr_group=16 cifhr_warps=2,cifhr_group=16
bash scripts/gpu_tune_sweep.sh ${TAG}_crowd crowd_1281_50p default cifhr_warps=3 cifhr_warps=8 cifhr_warps=6
bash scripts/gpu_tune_sweep.sh ${TAG}_dense dense_641_5p default cifhr_warps=2 cifhr_warps=3
