#!/bin/bash
# GPU parity tests, smoke() and one default bench line (summary printed)
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py --steps 20 2>/dev/null | tail -1 > gpurun_out/sanity.json
python -c "
import json
d=json.loads(open('gpurun_out/sanity.json').read())
print('value %.0f e2e %.0f full copy %.0f splat frac %.3f cpu %.0f' % (d['value'], d['e2e']['value'], d['e2e_full_copy']['value'], d['roofline']['frac'], d['cpu_baseline']['value']))
"
