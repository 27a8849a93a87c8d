#!/bin/bash
# value against the number of batches in flight.  $1 tag, $2 workload, rest: lane counts
OUT=gpurun_out; TAG=${1:-lanes}; WL=${2:-coco17_641_5p}; shift; shift
for l in "$@"; do
  python bench.py --value-only --no-cpu-baseline --distinct 2 --workload $WL --lanes $l > $OUT/${TAG}_last.out 2> $OUT/${TAG}_last.err || tail -5 $OUT/${TAG}_last.err
  python -c "
import json
d=json.loads(open('$OUT/${TAG}_last.out').read().strip().splitlines()[-1])
print('$WL lanes %2d value %7.0f  step %.4f  one-lane %.4f' % ($l, d['value'], d['ms_per_step'], d['one_lane_ms_per_step']))"
done | tee -a $OUT/${TAG}_lanes.log
