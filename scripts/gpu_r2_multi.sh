#!/bin/bash
# multi-GPU measurements on one 8-GPU box: shard_4096 (strong scaling) and the h2d floor at N = 1, 2, 4, 8,
# the default bench at N = 8 (weak scaling, e2e) -- TAG=$1
TAG=${1:-r2s}; OUT=gpurun_out; mkdir -p $OUT
run() {  # N, output, command...
  local n=$1 out=$2; shift; shift
  if [ $n -eq 1 ]; then python "$@" > $out 2> ${out%.json}.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) "$@" > $out 2> ${out%.json}.err; fi
  echo "N=$n $(basename $out) rc=$? $(tail -c 300 $out | head -c 300)"
}
for n in 1 2 4 8; do run $n $OUT/${TAG}_h2d_floor_${n}gpu.json scripts/h2d_floor.py; done
for n in 1 2 4 8; do run $n $OUT/${TAG}_shard4096_${n}gpu.json bench.py --gpus $n --workload shard_4096 --no-cpu-baseline; done
run 8 $OUT/${TAG}_bench_8gpu.json bench.py --gpus 8 --no-cpu-baseline
run 1 $OUT/${TAG}_bench_1gpu.json bench.py --gpus 1 --no-cpu-baseline
python - <<PY
import json
def load(p):
    try: return json.loads(open(p).read().strip().splitlines()[-1])
    except Exception as e: return None
for n in (1,2,4,8):
    f=load('$OUT/${TAG}_h2d_floor_%dgpu.json'%n); s=load('$OUT/${TAG}_shard4096_%dgpu.json'%n)
    if f: print('floor N=%d strided %.1f GB/s (%.0f img/s) plain %.1f full %.1f GB/s'%(n,f['strided']['aggregate_gbs'],f['strided']['images_per_s_floor'],f['plain']['aggregate_gbs'],f['full']['aggregate_gbs']))
    if s: print('shard N=%d value %.0f e2e %.0f job %.4f / %.4f s breakdown %s / %s'%(n,s['value'],s['e2e']['value'],s['job_s']['median'],s['e2e']['job_s']['median'],s['breakdown'],s['e2e']['breakdown']))
for n in (1,8):
    d=load('$OUT/${TAG}_bench_%dgpu.json'%n)
    if d: print('bench N=%d value %.0f e2e %.0f dev %.0f copy %.0f breakdown %s'%(n,d['value'],d['e2e']['value'],d['e2e_device_fields']['value'],d['e2e_full_copy']['value'],d['e2e']['breakdown']))
PY
