#!/bin/bash
# value against lanes and CUDA_DEVICE_MAX_CONNECTIONS (hardware work queues), default and crowd workloads
for w in coco17_641_5p crowd_1281_50p; do
for conn in 8 32; do
for lanes in 8 12 16 24; do
  echo "== $w conn=$conn lanes=$lanes"
  CUDA_DEVICE_MAX_CONNECTIONS=$conn python bench.py --workload $w --no-cpu-baseline --steps 24 --lanes $lanes 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('value %.0f one_lane %.0f' % (d['value'], d['value_one_lane']['value']))
    elif 'rror' in l: print(l.strip()[:200])
"
done; done; done
