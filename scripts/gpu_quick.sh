#!/bin/bash
# quick GPU check: parity tests + bench lines; extra env settings can be swept via $2 (e.g. "PP_CIFHR_CTAS=3 PP_CIFHR_CTAS=5")
TAG=${1:-q}
OUT=gpurun_out
python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; tail -3 $OUT/${TAG}_pytest.log
for setting in "" $2; do
echo "== $setting"
env $setting python bench.py --no-cpu-baseline --steps 20 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value %.0f e2e %.0f splat frac %.3f'%(d['value'],d['e2e']['value'],d['roofline']['frac'])); print(d['stages_ms'])
    elif 'rror' in l: print(l.strip()[:300])
"
done
