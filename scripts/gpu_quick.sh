#!/bin/bash
# quick GPU check: parity tests + one bench line
TAG=${1:-q}
OUT=gpurun_out
python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; tail -3 $OUT/${TAG}_pytest.log
python bench.py --no-cpu-baseline --steps 20 $2 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value %.0f e2e %.0f splat frac %.3f'%(d['value'],d['e2e']['value'],d['roofline']['frac'])); print(d['stages_ms'])
    elif 'rror' in l: print(l.strip())
"
for wl in crowd_1281_50p dense_641_5p; do
python bench.py --no-cpu-baseline --steps 10 --workload $wl 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$wl value %.0f e2e %.0f splat frac %.3f'%(d['value'],d['e2e']['value'],d['roofline']['frac'])); print(d['stages_ms'])
    elif 'rror' in l: print(l.strip())
"
done
