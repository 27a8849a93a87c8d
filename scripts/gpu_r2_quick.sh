#!/bin/bash
# quick check on the GPU box: parity tests, then the default / crowd / dense bench lines without the CPU baseline
TAG=${1:-q}; OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $OUT/${TAG}_pytest.log
for wl in coco17_641_5p crowd_1281_50p dense_641_5p; do
  steps=20; [ $wl != coco17_641_5p ] && steps=32
  timeout 600 python bench.py --workload $wl --steps $steps --no-cpu-baseline > $OUT/${TAG}_$wl.json 2> $OUT/${TAG}_$wl.err
  python - <<PY
import json
try:
    d=json.loads(open('$OUT/${TAG}_$wl.json').read().strip().splitlines()[-1])
    print('$wl value %.0f one_lane %.0f e2e %.0f sync %.0f frac %.3f lat %s'%(d['value'],d['value_one_lane']['value'],d['e2e']['value'],d['e2e']['one_call_per_step']['value'],d['roofline']['frac'],d['latency_one_image']))
    print(' stages', {k:round(v,4) for k,v in d['stages_ms'].items()})
except Exception as e:
    print('$wl failed', e); print(open('$OUT/${TAG}_$wl.err').read()[-1500:])
PY
done
