#!/bin/bash
TAG=${1:-r2u}; OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 $OUT/${TAG}_pytest.log
bash scripts/gpu_tune_sweep.sh ${TAG}_default coco17_641_5p default cifhr_warps=2
python scripts/determinism_stress.py 10 2>&1 | tail -4
