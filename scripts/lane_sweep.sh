#!/bin/bash
# sweep lanes in flight (throughput mode); prints value / one-lane / in-flight stage times
for lanes in ${1:-4 8 12 16}; do
  echo "== lanes=$lanes"
  python bench.py --no-cpu-baseline --steps 20 --lanes $lanes 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value %.0f one_lane %.0f e2e %.0f full copy %.0f'%(d['value'],d['value_one_lane']['value'],d['e2e']['value'],d['e2e_full_copy']['value'])); print({k:round(v,3) for k,v in d['stages_ms_in_flight'].items()})
    elif 'rror' in l: print(l.strip()[:300])
"
done
