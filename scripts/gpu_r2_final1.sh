#!/bin/bash
# final single-GPU set: parity suite, bench lines of every workload, reference arm, shard_4096, h2d floor
TAG=${1:-r3}; OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/${TAG}_pytest_gpu.log
timeout 600 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > $OUT/${TAG}_bench_ref.json 2> $OUT/${TAG}_bench_ref.err; echo "ref rc=$?"
: > $OUT/${TAG}_workloads.jsonl
for wl in ${WLS:-crowd_1281_50p dense_641_5p dense_641_5p_b256 cifdet_641_20b randinit_641_tamed randinit_641}; do
  steps="--steps 32"; case $wl in randinit*) steps="";; esac
  timeout 1200 python bench.py --workload $wl $steps >> $OUT/${TAG}_workloads.jsonl 2> $OUT/${TAG}_$wl.err; rc=$?; echo "$wl rc=$rc"
  [ $rc -ne 0 ] && tail -5 $OUT/${TAG}_$wl.err
done
timeout 900 python bench.py --workload shard_4096 > $OUT/${TAG}_shard4096_1gpu.json 2> $OUT/${TAG}_shard4096_1gpu.err; echo "shard rc=$?"
timeout 300 python scripts/h2d_floor.py > $OUT/${TAG}_h2d_floor_1gpu.json 2> $OUT/${TAG}_h2d_floor.err
timeout 300 python scripts/bench_head.py > $OUT/${TAG}_head.json 2> $OUT/${TAG}_head.err; cat $OUT/${TAG}_head.json
python - <<PY
import json
for line in open('$OUT/${TAG}_workloads.jsonl'):
    try:
        d=json.loads(line)
        print(d['config']['workload'][:28], 'value %.0f one-lane %.0f e2e %.0f dev %.0f cpu %s' % (d['value'], d.get('value_one_lane',{}).get('value',0), d.get('e2e',{}).get('value',0), d.get('e2e_device_fields',{}).get('value',0), d.get('cpu_baseline',{}).get('value')))
    except Exception as e: print('bad line', e)
PY
du -sh $OUT
