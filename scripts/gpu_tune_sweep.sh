#!/bin/bash
# value under different schedule knobs (pp_params tuning fields); $1 = tag, $2 = workload, rest = "knob=v,knob=v" sets
OUT=gpurun_out; TAG=${1:-tune}; WL=${2:-coco17_641_5p}; shift; shift
for t in "$@"; do
  arg=""; [ "$t" != "default" ] && arg="--tuning $t"
  python bench.py --value-only --no-cpu-baseline --workload $WL --distinct 2 $arg > $OUT/${TAG}_last.out 2> $OUT/${TAG}_last.err
  rc=$?
  if ! grep -q '"value"' $OUT/${TAG}_last.out; then
    echo "$t: FAILED rc=$rc"; tail -20 $OUT/${TAG}_last.err; cp $OUT/${TAG}_last.err "$OUT/${TAG}_fail_$t.err"; continue
  fi
  python -c "
import json,sys
d=json.loads(open('$OUT/${TAG}_last.out').read().strip().splitlines()[-1])
print('%-40s value %7.0f  step %.4f  one-lane %.4f  stages %s' % ('$t', d['value'], d['ms_per_step'], d['one_lane_ms_per_step'], d['stages_ms']))"
done | tee $OUT/${TAG}_sweep.log
