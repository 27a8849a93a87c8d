#!/bin/bash
# other workloads of BASELINE.json configs (parity cases / extra bench lines)
OUT=gpurun_out; TAG=${1:-w}
for wl in crowd_1281_50p dense_641_5p; do
  python bench.py --workload $wl --steps 10 > $OUT/${TAG}_$wl.json 2> $OUT/${TAG}_$wl.err
  python - <<PY
import json
try:
    d=json.loads(open('$OUT/${TAG}_$wl.json').read().strip().splitlines()[-1])
    print('$wl value %.0f e2e %.0f frac %.3f cpu %s'%(d['value'],d['e2e']['value'],d['roofline']['frac'],d.get('cpu_baseline',{}).get('value')))
    print(d['stages_ms'])
except Exception as e:
    print('$wl failed', e); print(open('$OUT/${TAG}_$wl.err').read()[-1500:])
PY
done
