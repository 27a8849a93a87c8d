#!/bin/bash
# quick GPU check: parity tests + short bench lines for a few tuning settings
TAG=${1:-q}
OUT=gpurun_out
python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; tail -3 $OUT/${TAG}_pytest.log
for hc in 4 8 16 64; do
  echo "== host_chunk $hc"
  python bench.py --no-cpu-baseline --steps 20 --host-chunk $hc 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value %.0f e2e %.0f splat %.4f ms frac %.3f'%(d['value'],d['e2e']['value'],d['stages_ms']['cifhr_splat'],d['roofline']['frac']))
    elif 'rror' in l: print(l.strip())
"
done
