#!/bin/bash
# stage-skip decomposition of the in-flight step (library rebuilt with PP_DEBUG_SKIP on the box)
OUT=gpurun_out; TAG=${1:-s}
PP_EXTRA_NVCC="-DPP_DEBUG_SKIP" python -c "from openpifpaf_b200 import build; build.build(force=True)" > $OUT/${TAG}_build.log 2>&1 || { tail $OUT/${TAG}_build.log; exit 1; }
python scripts/stage_skip_probe.py 2>&1 | tail -14 | tee $OUT/${TAG}_skip_probe.log
