#!/bin/bash
# Run on the GPU box (under gpurun): parity tests, bench, ncu launch list and one
# full capture of the CifHr splat kernel.  Outputs land in gpurun_out/.
set -u
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest_gpu.log 2>&1
echo "pytest rc=$?" >> $OUT/${TAG}_pytest_gpu.log
python bench.py --impl reference --steps 5 --warmup 1 > $OUT/${TAG}_bench_ref.json 2> $OUT/${TAG}_bench_ref.err
python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
echo "bench rc=$?" >> $OUT/${TAG}_bench.err
SHORT="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$SHORT > $OUT/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file $OUT/${TAG}_launches.csv $SHORT > $OUT/${TAG}_ncu_launches.log 2>&1
$SHORT > $OUT/${TAG}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:cifhr_splat -s 3 -c 2 \
    -f -o $OUT/${TAG}_splat $SHORT > $OUT/${TAG}_ncu_splat.log 2>&1
ls -la $OUT
