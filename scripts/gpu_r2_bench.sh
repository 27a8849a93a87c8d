#!/bin/bash
# round-2 measurement set on one GPU: parity tests, default bench + reference arm (two warm-ups, to look at the
# CPU baseline's spread), secondary workloads, shard_4096, h2d floor.  TAG=$1
TAG=${1:-r2}; OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/${TAG}_pytest.log
timeout 600 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"; tail -3 $OUT/${TAG}_bench.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > $OUT/${TAG}_bench_ref.json 2> $OUT/${TAG}_bench_ref.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 1 > $OUT/${TAG}_bench_ref_w1.json 2>/dev/null
timeout 900 python bench.py --workload shard_4096 > $OUT/${TAG}_shard4096_1gpu.json 2> $OUT/${TAG}_shard4096_1gpu.err; echo "shard rc=$?"; tail -3 $OUT/${TAG}_shard4096_1gpu.err
timeout 300 python scripts/h2d_floor.py > $OUT/${TAG}_h2d_floor_1gpu.json 2> $OUT/${TAG}_h2d_floor.err
python - <<PY
import json
def load(p):
    try: return json.loads(open(p).read().strip().splitlines()[-1])
    except Exception as e: return {'error': repr(e)}
d=load('$OUT/${TAG}_bench.json')
if 'value' in d:
    print('value %.0f (passes %d, %.2fs) one_lane %.0f e2e %.0f dev_fields %.0f sync %.0f copy %.0f frac %.3f hbm_step %.2f'%(d['value'],d['passes'],d['timed_region_s'],d['value_one_lane']['value'],d['e2e']['value'],d['e2e_device_fields']['value'],d['e2e']['one_call_per_step']['value'],d['e2e_full_copy']['value'],d['roofline']['frac'],d['hbm_step']['frac_of_peak']))
    print(' stages', {k:round(v,4) for k,v in d['stages_ms'].items()}); print(' lat', d['latency_one_image'].get('cuda_tensors_ms'), d['latency_one_image'].get('numpy_arrays_ms'))
    print(' e2e breakdown', d['e2e']['breakdown'], d['e2e_device_fields']['breakdown'], 'd2h', d['e2e']['d2h_bytes_per_step'])
    print(' cpu', {k:v for k,v in d.get('cpu_baseline',{}).items() if k!='sample'}); print(' clocks', d['clocks'])
else: print(d)
for n in ('bench_ref','bench_ref_w1'):
    r=load('$OUT/${TAG}_%s.json'%n); print(n, r.get('value'), {k:v for k,v in r.get('cpu_baseline',{}).items() if k!='sample'})
s=load('$OUT/${TAG}_shard4096_1gpu.json')
if 'value' in s: print('shard value %.0f e2e %.0f job_s %s / %s breakdown %s / %s checked %s gen %.1fs'%(s['value'],s['e2e']['value'],s['job_s'],s['e2e']['job_s'],s['breakdown'],s['e2e']['breakdown'],s['oracle_checked_images'],s['field_generation_s']))
else: print(s)
print(load('$OUT/${TAG}_h2d_floor_1gpu.json'))
PY
