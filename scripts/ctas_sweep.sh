#!/bin/bash
# value and stand-alone splat time against the number of persistent splat CTAs per SM
for c in 1 2 3 4; do
  echo "== PP_CIFHR_CTAS=$c"
  PP_CIFHR_CTAS=$c python bench.py --no-cpu-baseline --steps 24 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value %.0f one_lane %.0f splat %.4f ms in flight %.3f' % (d['value'], d['value_one_lane']['value'], d['stages_ms']['cifhr_splat'], d['stages_ms_in_flight']['cifhr_splat']))
"
done
