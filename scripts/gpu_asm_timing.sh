#!/bin/bash
# parity tests with the shipped build, then the phase breakdown of assemble_kernel for PP_ASM_WARPS in $2..
# (rebuilt on the box with the phase timer; the box is scratch)
OUT=gpurun_out; TAG=${1:-t}; shift
timeout 600 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/${TAG}_pytest.log
for W in ${@:-8}; do
  echo "== PP_ASM_WARPS=$W"
  PP_EXTRA_NVCC="-DPP_ASM_TIMING -DPP_ASM_WARPS=$W" python -c "from openpifpaf_b200 import build; build.build(force=True)" > $OUT/${TAG}_build.log 2>&1 || { tail $OUT/${TAG}_build.log; exit 1; }
  for wl in coco17_641_5p crowd_1281_50p dense_641_5p; do python scripts/asm_timing.py $wl 2>&1 | tail -11; done
  python bench.py --no-cpu-baseline --steps 20 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('bench (timer build) value %.0f one_lane %.0f assemble %.4f'%(d['value'],d['value_one_lane']['value'],d['stages_ms']['assemble']))"
done 2>&1 | tee $OUT/${TAG}_asm_timing.log
