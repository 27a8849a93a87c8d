#!/bin/bash
# value under different schedule knobs (pp_params tuning fields); $1 = tag, rest = "knob=v,knob=v" sets
OUT=gpurun_out; TAG=${1:-tune}; shift
for t in "$@"; do
  arg=""; [ "$t" != "default" ] && arg="--tuning $t"
  python bench.py --value-only --no-cpu-baseline $arg 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('%-44s value %7.0f  step %.4f  one-lane %.4f  in-flight %s' % ('$t', d['value'], d['ms_per_step'], d['one_lane_ms_per_step'], d['stages_ms_in_flight']))"
done | tee $OUT/${TAG}_sweep.log
