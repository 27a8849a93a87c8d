#!/bin/bash
# compute-sanitizer is closed on this pool: (1) the whole GPU suite on a -DPP_DEVICE_ASSERT build (bounds
# checks in the kernels trap), (2) repeat-run determinism with 8 batches in flight.  Restores the
# production build afterwards (the box copy only).  TAG=$1
set -u
TAG=${1:-r2t}
OUT=gpurun_out
mkdir -p $OUT
python scripts/determinism_stress.py 40 > $OUT/${TAG}_determinism.log 2>&1
echo "rc=$?" >> $OUT/${TAG}_determinism.log
tail -4 $OUT/${TAG}_determinism.log
PP_EXTRA_NVCC=-DPP_DEVICE_ASSERT python -c "
from openpifpaf_b200 import build
print(build.build(force=True))" > $OUT/${TAG}_assert_build.log 2>&1
cuobjdump -elf openpifpaf_b200/libpifpaf_b200.so | grep -c __assertfail > $OUT/${TAG}_assert_sites.txt
timeout 1200 python -m pytest tests -m gpu -q > $OUT/${TAG}_assert_pytest.log 2>&1
echo "pytest rc=$?" >> $OUT/${TAG}_assert_pytest.log
tail -3 $OUT/${TAG}_assert_pytest.log
timeout 600 python scripts/sanitize_case.py > $OUT/${TAG}_assert_cases.log 2>&1
echo "rc=$?" >> $OUT/${TAG}_assert_cases.log
tail -3 $OUT/${TAG}_assert_cases.log
