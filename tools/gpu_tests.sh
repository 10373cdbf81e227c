#!/bin/bash
# GPU test suite (usage under gpurun: bash tools/gpu_tests.sh <tag>): the whole suite, then the no-alloc test alone in a fresh process
set -u
OUT=gpurun_out/${1:-tests}
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log
tail -40 $OUT/pytest_gpu.log | cut -c1-700
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $OUT/pytest_noalloc_alone.log 2>&1; echo "no-alloc alone rc=$?" | tee -a $OUT/pytest_noalloc_alone.log
tail -5 $OUT/pytest_noalloc_alone.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $OUT/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/smoke.log
